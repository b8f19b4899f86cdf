"""Batching HTTP front door for the reference's Gradio REST API (second half).

`client_streaming.sh` (reference, lines 52-116) talks to `gradio_voice_converter_unstreaming_streaming.py` in three calls:

    POST /gradio_api/upload                                     multipart `files=@input  files=@reference`  -> ["<path>", "<path>"]
    POST /gradio_api/call/process_audio_nonstreaming            {"data": [{"path": ...}, {"path": ...}, 0.8]}  -> {"event_id": "..."}
    (`process_audio_streaming` likewise, data = [input, reference, block_size, max_token_len]: chunked decode as stream_inference)
    GET  /gradio_api/call/process_audio_nonstreaming/<event_id> server-sent events: `event: heartbeat` while working, then
                                                                `event: complete` + `data: [html, info, file]` where html embeds
                                                                `data:audio/wav;base64,...` (the client's regex), or `event: error`

This module serves exactly those endpoints with aiohttp (gradio itself is not needed and not installed here) and puts the
decoder behind them.  The reference handler decodes one request at a time (`process_audio_nonstreaming`, lines 206-289:
tokenize input, `model.decode([tokens], prompt_speech=reference)`); here every request that arrives while a batch is being
assembled joins it, and the batch goes through ONE `AudioDecoder.decode_batch` call -- the config-2 path.

Scope: turning audio into speech tokens (and the reference audio into prompt tokens / x-vector) is the tokenizer side,
outside this repository (SURVEY.md §8 "out of scope").  The two steps are therefore hooks:

    tokenize(path)  -> 1-D integer tokens of the utterance to synthesise
    reference(path) -> (prompt_token [1, P] int, prompt_feat [1, 4 P, 80] float, embedding [1, 192] float)

The default hooks accept PRE-TOKENISED uploads so the front door is usable and testable on its own: the input file is a
`.npy` / `.json` of token ids, the reference file an `.npz` with `prompt_token`, `embedding` and either `prompt_feat` or
`prompt_wav` (24 kHz mono, run through the decoder's own `feat_extractor`, whisper_encoder_decoder.py:191-195).  A deployment
that has the reference's tokenizer passes `tokenize=model.encode_token` and a `reference` built on its frontend.
"""
import asyncio
import base64
import io
import json
import os
import tempfile
import time
import uuid
import wave

import numpy as np

API = "process_audio_nonstreaming"
API_STREAMING = "process_audio_streaming"     # same calls, data = [input, reference, block_size, max_token_len]
SAMPLE_RATE = 24000


def wav_bytes(samples, sample_rate=SAMPLE_RATE):
    """float waveform in [-1, 1] -> 16-bit PCM WAV file contents (what save_audio_for_gradio writes, 24 kHz mono)."""
    pcm = (np.clip(np.asarray(samples, np.float32).reshape(-1), -1.0, 1.0) * 32767.0).astype("<i2")
    buf = io.BytesIO()
    with wave.open(buf, "wb") as w:
        w.setnchannels(1)
        w.setsampwidth(2)
        w.setframerate(sample_rate)
        w.writeframes(pcm.tobytes())
    return buf.getvalue()


def audio_html(wav, filename):
    """The HTML the reference returns (get_audio_html, gradio_voice_converter_unstreaming_streaming.py:180-200): an <audio> tag
    and a download link, both carrying the file as a base64 data URL."""
    b64 = base64.b64encode(wav).decode("ascii")
    return (f'<div><audio controls><source src="data:audio/wav;base64,{b64}" type="audio/wav"></audio>'
            f'<a href="data:audio/wav;base64,{b64}" download="{filename}" target="_blank">download</a></div>')


def default_tokenize(path):
    """Pre-tokenised input: `.npy` or `.json` holding the utterance's speech token ids."""
    import torch
    if path.endswith(".json"):
        toks = np.asarray(json.load(open(path)), np.int64)
    elif path.endswith(".npy"):
        toks = np.load(path).astype(np.int64)
    else:
        raise ValueError("no tokenizer attached: upload the utterance as .npy / .json token ids, or construct FrontDoor with "
                         "tokenize=<audio file -> tokens> (the reference's encode_token; outside this repository's scope)")
    if toks.ndim != 1 or toks.size == 0:
        raise ValueError("token file must hold a non-empty 1-D array of token ids")
    return torch.from_numpy(toks)


def make_default_reference(decoder):
    def reference(path):
        """Pre-processed speaker reference: `.npz` with prompt_token [P], embedding [192] and prompt_feat [F, 80] or prompt_wav [L]."""
        import torch
        if not path.endswith(".npz"):
            raise ValueError("no reference frontend attached: upload the reference as .npz (prompt_token, embedding, prompt_feat | "
                             "prompt_wav), or construct FrontDoor with reference=<audio file -> (prompt_token, prompt_feat, embedding)>")
        z = np.load(path)
        ptok = torch.from_numpy(z["prompt_token"].astype(np.int64)).reshape(1, -1)
        emb = torch.from_numpy(z["embedding"].astype(np.float32)).reshape(1, -1)
        if "prompt_feat" in z.files:
            feat = torch.from_numpy(z["prompt_feat"].astype(np.float32)).reshape(1, -1, 80)
        else:   # whisper_encoder_decoder.py:191-195: mel of the 24 kHz prompt through the decoder's own feat_extractor
            wav = torch.from_numpy(z["prompt_wav"].astype(np.float32)).reshape(1, -1)
            feat = decoder.feat_extractor(wav.to(decoder.device)).squeeze(0).transpose(0, 1).float().cpu().unsqueeze(0)
        n = min(feat.shape[1] // 4, ptok.shape[1])      # whisper_encoder_decoder.py:224-227: 4 mel frames per prompt token
        return ptok[:, :n], feat[:, : 4 * n], emb
    return reference


class _Job:
    def __init__(self, tokens, voice, stream_params=None):
        self.tokens, self.voice = tokens, voice
        self.stream_params = stream_params      # None: one-shot decode; (block_size, max_token_len): chunked like stream_inference
        self.done = asyncio.Event()
        self.result = None      # (html, info, file dict)
        self.error = None
        self.t0 = time.time()


class FrontDoor:
    """The three Gradio REST endpoints over one batching queue.  `decoder` needs `decode_batch(tokens, prompt_tokens,
    prompt_feats, embeddings)` returning {"wav": ragged waveform, "wav_lens": [...]} (flow_inference.AudioDecoder)."""

    def __init__(self, decoder, tokenize=None, reference=None, max_batch=64, window_s=0.005, heartbeat_s=1.0, upload_dir=None):
        self.decoder = decoder
        self.tokenize = tokenize or default_tokenize
        self.reference = reference or make_default_reference(decoder)
        self.max_batch, self.window_s, self.heartbeat_s = max_batch, window_s, heartbeat_s
        self.upload_dir = upload_dir or tempfile.mkdtemp(prefix="mossdec_upload_")
        self.jobs = {}
        self.queue = None
        self._task = None
        self.batches = []        # sizes of the decode_batch calls made (observability / tests)

    # ---- batching: everything queued when a batch closes goes through one decode_batch call
    async def _batch_loop(self):
        loop = asyncio.get_running_loop()
        while True:
            first = await self.queue.get()
            batch = [first]
            deadline = loop.time() + self.window_s
            while len(batch) < self.max_batch:
                timeout = deadline - loop.time()
                if timeout <= 0:
                    break
                try:
                    batch.append(await asyncio.wait_for(self.queue.get(), timeout))
                except asyncio.TimeoutError:
                    break
            while len(batch) < self.max_batch and not self.queue.empty():
                batch.append(self.queue.get_nowait())
            # one decode per kind: one-shot requests together, streaming requests of equal (block_size, max_token_len) together
            rest = [j for j in batch if j.stream_params != first.stream_params]
            batch = [j for j in batch if j.stream_params == first.stream_params]
            for j in rest:
                self.queue.put_nowait(j)
            self.batches.append(len(batch))
            try:
                wavs = await loop.run_in_executor(None, self._decode, batch)
                for job, wav in zip(batch, wavs):
                    data = wav_bytes(wav)
                    name = f"nonstreaming_output_{uuid.uuid4().hex[:8]}.wav"
                    path = os.path.join(self.upload_dir, name)
                    with open(path, "wb") as f:
                        f.write(data)
                    n = int(np.asarray(wav).size)
                    info = (f"decoded {len(job.tokens)} tokens -> {n / SAMPLE_RATE:.2f} s of audio in a batch of {len(batch)} "
                            f"({time.time() - job.t0:.3f} s after the request)")
                    job.result = (audio_html(data, name), info, {"path": path, "orig_name": name, "meta": {"_type": "gradio.FileData"}})
            except Exception as e:   # noqa: BLE001 - reported to every client of the batch as an error event
                for job in batch:
                    job.error = f"{type(e).__name__}: {e}"
            for job in batch:
                job.done.set()

    def _decode(self, batch):
        import torch
        if batch[0].stream_params is not None:
            # process_audio_streaming (gradio_voice_converter_unstreaming_streaming.py:299-401 -> decode_streaming ->
            # stream_inference): chunked decode of every request of the batch through ONE batched stream scheduler
            from .streaming import StreamScheduler
            block, mtl = batch[0].stream_params
            sched = StreamScheduler(self.decoder, block, mtl)
            for j in batch:
                sched.add(j.tokens.reshape(-1).tolist(), *j.voice)
            return [w.detach().float().cpu().reshape(-1).numpy() for w in sched.run()]
        toks = [j.tokens.reshape(1, -1) for j in batch]
        ptoks = [j.voice[0] for j in batch]
        pfeats = [j.voice[1] for j in batch]
        embs = torch.cat([j.voice[2].reshape(1, -1) for j in batch], 0)
        out = self.decoder.decode_batch(toks, ptoks, pfeats, embs)
        # AudioDecoder.decode_batch: ragged waveform {"wav": [sum 1920 N_b], "wav_lens": [...]}; one D2H copy for the batch
        wav = out["wav"].detach().float().cpu().numpy().reshape(-1)
        wavs, off = [], 0
        for n in out["wav_lens"]:
            wavs.append(wav[off:off + int(n)])
            off += int(n)
        return wavs

    def _ensure_loop(self):
        if self.queue is None:
            self.queue = asyncio.Queue()
        if self._task is None or self._task.done():
            self._task = asyncio.get_running_loop().create_task(self._batch_loop())

    # ---- POST /gradio_api/upload
    async def upload(self, request):
        from aiohttp import web
        paths = []
        reader = await request.multipart()
        while True:
            part = await reader.next()
            if part is None:
                break
            if part.name != "files":
                continue
            name = os.path.basename(part.filename or "upload.bin")
            path = os.path.join(self.upload_dir, f"{uuid.uuid4().hex[:12]}_{name}")
            with open(path, "wb") as f:
                while True:
                    chunk = await part.read_chunk()
                    if not chunk:
                        break
                    f.write(chunk)
            paths.append(path)
        return web.json_response(paths)

    def _path_of(self, item):
        path = item.get("path") if isinstance(item, dict) else item
        if not isinstance(path, str) or os.path.dirname(os.path.abspath(path)) != os.path.abspath(self.upload_dir):
            raise ValueError("data[i].path must be a path returned by /gradio_api/upload")
        return path

    # ---- POST /gradio_api/call/process_audio_nonstreaming
    async def call(self, request):
        from aiohttp import web
        try:
            body = await request.json()
            data = body["data"]
            tokens = self.tokenize(self._path_of(data[0]))
            voice = self.reference(self._path_of(data[1]))
            stream_params = None
            if request.match_info.get("api") == API_STREAMING:
                block = int(data[2])
                mtl = None if len(data) < 4 or data[3] in (None, 0, "") else int(data[3])
                if block < 1 or (mtl is not None and mtl < block + 3):
                    raise ValueError("block_size must be >= 1 and max_token_len >= block_size + 3 (the look-ahead)")
                stream_params = (block, mtl)
        except Exception as e:   # noqa: BLE001 - the client prints the body of a failed request
            return web.json_response({"error": f"{type(e).__name__}: {e}"}, status=422)
        self._ensure_loop()
        job = _Job(tokens, voice, stream_params)
        event_id = uuid.uuid4().hex
        self.jobs[event_id] = job
        self.queue.put_nowait(job)
        return web.json_response({"event_id": event_id})

    # ---- GET /gradio_api/call/process_audio_nonstreaming/<event_id>   (server-sent events)
    async def events(self, request):
        from aiohttp import web
        job = self.jobs.get(request.match_info["event_id"])
        if job is None:
            return web.json_response({"error": "unknown event_id"}, status=404)
        resp = web.StreamResponse(headers={"Content-Type": "text/event-stream", "Cache-Control": "no-cache"})
        await resp.prepare(request)
        while not job.done.is_set():
            try:
                await asyncio.wait_for(job.done.wait(), self.heartbeat_s)
            except asyncio.TimeoutError:
                await resp.write(b"event: heartbeat\ndata: null\n\n")
        if job.error is not None:
            await resp.write(b"event: error\ndata: " + json.dumps(job.error).encode() + b"\n\n")
        else:
            await resp.write(b"event: complete\ndata: " + json.dumps(list(job.result)).encode() + b"\n\n")
        await resp.write_eof()
        self.jobs.pop(request.match_info["event_id"], None)
        return resp


def make_app(door):
    from aiohttp import web
    app = web.Application(client_max_size=256 << 20)
    app.router.add_post("/gradio_api/upload", door.upload)
    app.router.add_post(f"/gradio_api/call/{{api:{API}|{API_STREAMING}}}", door.call)
    app.router.add_get(f"/gradio_api/call/{{api:{API}|{API_STREAMING}}}/{{event_id}}", door.events)
    return app


def main(argv=None):
    import argparse
    from aiohttp import web
    from .flow_inference import AudioDecoder
    ap = argparse.ArgumentParser(description="batching front door for the reference's Gradio REST API (client_streaming.sh)")
    ap.add_argument("--config", required=True)
    ap.add_argument("--flow-ckpt", required=True)
    ap.add_argument("--hift-ckpt", required=True)
    ap.add_argument("--campplus", default=None)
    ap.add_argument("--port", type=int, default=7864)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--window-ms", type=float, default=5.0)
    a = ap.parse_args(argv)
    dec = AudioDecoder(a.config, a.flow_ckpt, a.hift_ckpt, a.campplus, device="cuda")
    web.run_app(make_app(FrontDoor(dec, max_batch=a.max_batch, window_s=a.window_ms / 1e3)), port=a.port)


if __name__ == "__main__":
    main()

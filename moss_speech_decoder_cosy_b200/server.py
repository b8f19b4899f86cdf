"""Decoder-backed transport, the wire protocol of the reference's server.py.

The reference server (server.py:9-99) is a WebSocket skeleton on `/api/chat`: binary messages whose first byte is the kind
(0 handshake, 1 audio, 2 text), Opus both ways through `sphn`, 1920-sample frames at 24 kHz, one connection at a time behind
an asyncio lock, and the decoder call commented out (server.py:69-75 echoes the microphone back).  This module keeps that
protocol and puts the decoder behind it:

  * kind 1 (audio, client -> server) is decoded and dropped -- the speech language model that would turn it into tokens is
    outside this repository's scope (SURVEY.md §8 "out of scope");
  * kind 3 (client -> server) carries the speech tokens that language model produces: little-endian int32 codes;
  * kind 4 (client -> server) ends the utterance (the reference's `llm_end_dict[uuid] = True`, flow_inference.py:205);
  * kind 1 (audio, server -> client) carries the generated waveform in 1920-sample frames, kind 4 (server -> client) says the
    utterance is complete.

Instead of the reference's one-connection lock, every connection is a `Stream` of one shared `StreamScheduler`; one batching
task steps the scheduler whenever any stream has a decodable chunk, so concurrent connections share each estimator launch
(BASELINE config 3).  The scheduler step runs in a worker thread: the event loop keeps receiving tokens meanwhile.

Codec: `sphn` (Opus) when importable -- it is not in this image -- else 16-bit little-endian PCM; the handshake says which
(b"\\x00" alone = Opus as in the reference; b"\\x00pcm16" = PCM).
"""
import asyncio

import numpy as np

KIND_HANDSHAKE, KIND_AUDIO, KIND_TEXT, KIND_TOKENS, KIND_END = 0, 1, 2, 3, 4
FRAME_SIZE = 1920                      # server.py:14
SAMPLE_RATE = 24000                    # server.py:104


class Pcm16Codec:
    """Stand-in for sphn.OpusStreamWriter/Reader when sphn is absent: whole frames of int16 little-endian samples."""
    name = b"pcm16"

    def __init__(self, sample_rate):
        self.sample_rate = sample_rate
        self._out, self._in = [], []

    def append_pcm(self, frame):
        self._out.append((np.clip(np.asarray(frame, np.float32), -1.0, 1.0) * 32767.0).astype("<i2").tobytes())

    def read_bytes(self):
        out, self._out = b"".join(self._out), []
        return out

    def append_bytes(self, payload):
        self._in.append(np.frombuffer(payload[: len(payload) // 2 * 2], "<i2").astype(np.float32) / 32767.0)

    def read_pcm(self):
        if not self._in:
            return np.zeros(0, np.float32)
        out, self._in = np.concatenate(self._in), []
        return out


class OpusCodec:
    """sphn's Opus stream pair behind the same four calls (server.py:95-96)."""
    name = b""

    def __init__(self, sample_rate):
        import sphn
        self._w, self._r = sphn.OpusStreamWriter(sample_rate), sphn.OpusStreamReader(sample_rate)

    def append_pcm(self, frame):
        self._w.append_pcm(np.asarray(frame, np.float32))

    def read_bytes(self):
        return self._w.read_bytes()

    def append_bytes(self, payload):
        self._r.append_bytes(payload)

    def read_pcm(self):
        return self._r.read_pcm()


def default_codec():
    try:
        import sphn  # noqa: F401
        return OpusCodec
    except ImportError:
        return Pcm16Codec


class _Connection:
    def __init__(self, index, codec):
        self.index, self.codec = index, codec
        self.pending = np.zeros(0, np.float32)       # generated samples not yet cut into frames
        self.finished = asyncio.Event()


class ServerState:
    """server.py:9-99 with the decoder attached.  `scheduler` is a streaming.StreamScheduler (or anything with its
    `add(tokens, prompt_token, prompt_feat, embedding, ended)`, `streams`, `step()`); `voice` = (prompt_token, prompt_feat,
    embedding) of the speaker every connection uses."""

    def __init__(self, scheduler, voice, sample_rate=SAMPLE_RATE, codec=None, idle_sleep=0.001):
        self.scheduler, self.voice = scheduler, voice
        self.frame_size, self.sample_rate = FRAME_SIZE, sample_rate
        self.codec_cls = codec or default_codec()
        self.idle_sleep = idle_sleep
        self.conns = {}
        self._task = None
        self._lock = asyncio.Lock()                 # streams are added / released only between scheduler steps
        self.steps = 0
        self.max_batch = 0

    # ---- the batching loop: one for all connections ----
    def _ready(self):
        return any((not s.done) and s.next_chunk() is not None for s in self.scheduler.streams)

    async def _batch_loop(self):
        loop = asyncio.get_running_loop()
        while self.conns:
            if not self._ready():
                await asyncio.sleep(self.idle_sleep)
                continue
            async with self._lock:
                emitted = await loop.run_in_executor(None, self.scheduler.step)
            self.steps += 1
            self.max_batch = max(self.max_batch, len(emitted))
            for i, wav in emitted.items():
                conn = self.conns.get(i)
                if conn is None:
                    continue
                wav = np.asarray(wav.detach().float().cpu().reshape(-1) if hasattr(wav, "detach") else wav, np.float32).reshape(-1)
                conn.pending = np.concatenate((conn.pending, wav))
                last = self.scheduler.streams[i].done
                while conn.pending.shape[0] >= self.frame_size:
                    conn.codec.append_pcm(conn.pending[: self.frame_size])
                    conn.pending = conn.pending[self.frame_size:]
                if last:
                    if conn.pending.shape[0]:            # tail frame, zero padded to the frame size
                        tail = np.zeros(self.frame_size, np.float32)
                        tail[: conn.pending.shape[0]] = conn.pending
                        conn.codec.append_pcm(tail)
                        conn.pending = np.zeros(0, np.float32)
                    conn.finished.set()

    def _ensure_loop(self):
        if self._task is None or self._task.done():
            self._task = asyncio.get_running_loop().create_task(self._batch_loop())

    # ---- one connection ----
    async def handle_chat(self, request):
        import aiohttp
        from aiohttp import web
        ws = web.WebSocketResponse()
        await ws.prepare(request)
        codec = self.codec_cls(self.sample_rate)
        async with self._lock:
            index = self.scheduler.add([], *self.voice, ended=False)
            stream = self.scheduler.streams[index]
            conn = _Connection(index, codec)
            self.conns[index] = conn
        self._ensure_loop()
        close = False

        async def recv_loop():
            nonlocal close
            try:
                async for message in ws:
                    if message.type == aiohttp.WSMsgType.ERROR:
                        break
                    if message.type != aiohttp.WSMsgType.BINARY or len(message.data) == 0:
                        continue                          # server.py:31-40: ignored with a warning
                    kind, payload = message.data[0], message.data[1:]
                    if kind == KIND_AUDIO:
                        codec.append_bytes(payload)
                        codec.read_pcm()                  # microphone audio: no tokenizer in scope, dropped
                    elif kind == KIND_TOKENS and not stream.ended:
                        stream.push(np.frombuffer(payload[: len(payload) // 4 * 4], "<i4").tolist())
                    elif kind == KIND_END:
                        stream.finish()
            finally:
                close = True
                if not stream.ended:
                    stream.finish()

        async def send_loop():
            told = False
            while not close:
                msg = codec.read_bytes()
                if len(msg) > 0:
                    await ws.send_bytes(bytes([KIND_AUDIO]) + msg)
                elif conn.finished.is_set() and not told:
                    await ws.send_bytes(bytes([KIND_END]))
                    told = True
                else:
                    await asyncio.sleep(self.idle_sleep)

        try:
            await ws.send_bytes(bytes([KIND_HANDSHAKE]) + codec.name)     # server.py:98
            await asyncio.gather(recv_loop(), send_loop())
        finally:
            async with self._lock:
                self.conns.pop(index, None)
                if hasattr(self.scheduler, "release"):
                    self.scheduler.release(index)          # position and prefix-cache slot go to the next connection
                else:
                    stream.finish()
        return ws


def make_app(state):
    from aiohttp import web
    app = web.Application()
    app.router.add_get("/api/chat", state.handle_chat)               # server.py:106
    return app


def main(argv=None):
    import argparse
    import torch
    from aiohttp import web
    from .flow_inference import AudioDecoder
    from .streaming import StreamScheduler
    ap = argparse.ArgumentParser(description="token -> waveform WebSocket server (reference server.py protocol)")
    ap.add_argument("--config", required=True)
    ap.add_argument("--flow-ckpt", required=True)
    ap.add_argument("--hift-ckpt", required=True)
    ap.add_argument("--campplus", required=True)
    ap.add_argument("--prompt-wav", required=True, help="24 kHz mono float32 .npy of the speaker prompt (-> prompt mel)")
    ap.add_argument("--prompt-tokens", required=True, help=".npy of the prompt's speech tokens")
    ap.add_argument("--embedding", required=True, help=".npy of the speaker's 192-d campplus x-vector")
    ap.add_argument("--port", type=int, default=8023)
    ap.add_argument("--block-size", type=int, default=5)
    ap.add_argument("--max-token-len", type=int, default=40)
    a = ap.parse_args(argv)
    dec = AudioDecoder(a.config, a.flow_ckpt, a.hift_ckpt, a.campplus, device="cuda")
    prompt = torch.from_numpy(np.load(a.prompt_wav)).float().reshape(1, -1)
    # whisper_encoder_decoder.py:192: the prompt mel through the decoder's own feat_extractor
    feat = dec.feat_extractor(prompt).squeeze(dim=0).transpose(0, 1).unsqueeze(0).float().cpu()
    emb = torch.from_numpy(np.load(a.embedding)).float().reshape(1, -1)
    ptok = torch.from_numpy(np.load(a.prompt_tokens)).long().reshape(1, -1)
    sched = StreamScheduler(dec, a.block_size, a.max_token_len)
    web.run_app(make_app(ServerState(sched, (ptok, feat, emb))), port=a.port)


if __name__ == "__main__":
    main()

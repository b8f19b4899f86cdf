"""HTTP front door: the three Gradio REST calls of the reference's client_streaming.sh (upload, call,
server-sent events with the base64 WAV in HTML) with cross-request batching into ONE decode_batch call.  CPU only: the decoder
is a deterministic stand-in; everything else (multipart upload, hooks, queue, SSE framing, WAV encoding) is the shipped code."""
import asyncio
import base64
import io
import json
import re
import wave

import numpy as np
import pytest
import torch

from moss_speech_decoder_cosy_b200 import http_frontdoor as fd

aiohttp = pytest.importorskip("aiohttp")
from aiohttp import FormData  # noqa: E402
from aiohttp.test_utils import TestClient, TestServer  # noqa: E402


class FakeDecoder:
    """decode_batch with AudioDecoder's return contract: 1920 samples per token, valued token / 8192 (+ speaker offset)."""
    device = "cpu"

    def __init__(self, fail=False):
        self.calls, self.fail = [], fail

    def decode_batch(self, tokens, prompt_tokens, prompt_feats, embeddings):
        if self.fail:
            raise RuntimeError("decoder exploded")
        self.calls.append(len(tokens))
        assert embeddings.shape == (len(tokens), 192)
        wavs = []
        for t, pt, pf in zip(tokens, prompt_tokens, prompt_feats):
            assert t.dim() == 2 and pt.dim() == 2 and pf.shape[1] == 4 * pt.shape[1] and pf.shape[2] == 80
            wavs.append(torch.repeat_interleave(t.reshape(-1).float() / 8192.0, 1920))
        return {"wav": torch.cat(wavs), "wav_lens": [w.numel() for w in wavs]}


def _npy(a):
    b = io.BytesIO()
    np.save(b, a)
    return b.getvalue()


def _npz(**kw):
    b = io.BytesIO()
    np.savez(b, **kw)
    return b.getvalue()


async def _one_request(tc, tokens, ref_bytes, in_name="tokens.npy"):
    """client_streaming.sh: upload both files, call, follow the event stream until the data line."""
    form = FormData()
    form.add_field("files", _npy(tokens) if in_name.endswith(".npy") else b"RIFFxxxx", filename=in_name)
    form.add_field("files", ref_bytes, filename="ref.npz")
    r = await tc.post("/gradio_api/upload", data=form)
    paths = await r.json()
    assert r.status == 200 and len(paths) == 2
    meta = {"_type": "gradio.FileData"}
    r = await tc.post(f"/gradio_api/call/{fd.API}", json={"data": [{"path": paths[0], "meta": meta}, {"path": paths[1], "meta": meta}, 0.8]})
    body = await r.json()
    if r.status != 200:
        return r.status, body
    r = await tc.get(f"/gradio_api/call/{fd.API}/{body['event_id']}")
    assert r.headers["Content-Type"].startswith("text/event-stream")
    text = (await r.read()).decode()
    return 200, text


def _wav_from_events(text):
    assert "event: complete" in text
    line = [l for l in text.splitlines() if re.match(r"^data: \[.+\]$", l)][0]      # the client's own test for the final line
    html, info, fileinfo = json.loads(line[len("data: "):])
    m = re.search(r"data:audio/wav;base64,([A-Za-z0-9+/=]+)", html)                  # ... and its regex for the payload
    with wave.open(io.BytesIO(base64.b64decode(m.group(1)))) as w:
        assert (w.getnchannels(), w.getsampwidth(), w.getframerate()) == (1, 2, 24000)
        pcm = np.frombuffer(w.readframes(w.getnframes()), "<i2")
    assert fileinfo["path"].endswith(".wav") and "tokens" in info
    return pcm.astype(np.float32) / 32767.0


def _ref(P=7):
    return _npz(prompt_token=np.arange(P), embedding=np.ones(192, np.float32), prompt_feat=np.zeros((4 * P + 3, 80), np.float32))


def test_frontdoor_protocol_and_cross_request_batching():
    async def main():
        dec = FakeDecoder()
        door = fd.FrontDoor(dec, window_s=0.05)
        async with TestClient(TestServer(fd.make_app(door))) as tc:
            rng = np.random.default_rng(0)
            toks = [rng.integers(0, 8192, n) for n in (25, 60, 41, 33)]
            got = await asyncio.gather(*[_one_request(tc, t, _ref()) for t in toks])
        return dec, door, toks, got
    dec, door, toks, got = asyncio.run(main())
    for t, (status, text) in zip(toks, got):
        assert status == 200
        wav = _wav_from_events(text)
        want = np.repeat(t.astype(np.float32) / 8192.0, 1920)
        assert wav.shape == want.shape and np.abs(wav - want).max() < 1e-4
    assert sum(dec.calls) == 4 and max(dec.calls) > 1, dec.calls          # concurrent requests shared a decode_batch call
    assert door.batches == dec.calls and not door.jobs                    # every job handed out and forgotten


def test_frontdoor_errors_are_loud():
    async def main():
        out = {}
        door = fd.FrontDoor(FakeDecoder(), heartbeat_s=0.01)
        async with TestClient(TestServer(fd.make_app(door))) as tc:
            # an audio file without a tokenizer hook: refused at the call, with the reason
            out["wav"] = await _one_request(tc, np.arange(5), _ref(), in_name="input.wav")
            # a path that did not come from /upload
            r = await tc.post(f"/gradio_api/call/{fd.API}", json={"data": [{"path": "/etc/passwd"}, {"path": "/etc/passwd"}, 0.8]})
            out["path"] = (r.status, await r.json())
            r = await tc.get(f"/gradio_api/call/{fd.API}/nope")
            out["unknown"] = r.status
        door2 = fd.FrontDoor(FakeDecoder(fail=True))
        async with TestClient(TestServer(fd.make_app(door2))) as tc:
            out["fail"] = await _one_request(tc, np.arange(9), _ref())
        return out
    out = asyncio.run(main())
    assert out["wav"][0] == 422 and "no tokenizer attached" in out["wav"][1]["error"]
    assert out["path"][0] == 422 and "upload" in out["path"][1]["error"]
    assert out["unknown"] == 404
    assert out["fail"][0] == 200 and "event: error" in out["fail"][1] and "decoder exploded" in out["fail"][1]


def test_frontdoor_hooks_replace_the_default_parsers():
    """A deployment with the reference's tokenizer passes its own hooks: any file type is then accepted."""
    async def main():
        dec = FakeDecoder()
        voice = (torch.arange(4).reshape(1, -1), torch.zeros(1, 16, 80), torch.zeros(1, 192))
        door = fd.FrontDoor(dec, tokenize=lambda path: torch.arange(12), reference=lambda path: voice)
        async with TestClient(TestServer(fd.make_app(door))) as tc:
            form = FormData()
            form.add_field("files", b"RIFF....", filename="in.mp3")
            form.add_field("files", b"RIFF....", filename="ref.wav")
            paths = await (await tc.post("/gradio_api/upload", data=form)).json()
            r = await tc.post(f"/gradio_api/call/{fd.API}", json={"data": [{"path": paths[0]}, {"path": paths[1]}, 0.8]})
            ev = (await r.json())["event_id"]
            text = (await (await tc.get(f"/gradio_api/call/{fd.API}/{ev}")).read()).decode()
        return text
    wav = _wav_from_events(asyncio.run(main()))
    assert wav.shape == (12 * 1920,)


def test_frontdoor_streaming_api_goes_through_the_stream_scheduler(monkeypatch):
    """process_audio_streaming: same three calls with [input, reference, block_size, max_token_len]; requests of equal
    parameters share one batched stream scheduler, one-shot requests queued at the same time are decoded separately."""
    from moss_speech_decoder_cosy_b200 import streaming
    made = []

    class FakeSched:
        def __init__(self, dec, block, mtl):
            self.args, self.items = (block, mtl), []
            made.append(self)

        def add(self, tokens, ptok, pfeat, emb):
            self.items.append(tokens)

        def run(self):
            return [torch.repeat_interleave(torch.tensor(t).float() / 8192.0, 1920).reshape(1, -1) for t in self.items]
    monkeypatch.setattr(streaming, "StreamScheduler", FakeSched)

    async def call(tc, api, tokens, extra):
        form = FormData()
        form.add_field("files", _npy(tokens), filename="tokens.npy")
        form.add_field("files", _ref(), filename="ref.npz")
        paths = await (await tc.post("/gradio_api/upload", data=form)).json()
        r = await tc.post(f"/gradio_api/call/{api}", json={"data": [{"path": paths[0]}, {"path": paths[1]}] + extra})
        body = await r.json()
        if r.status != 200:
            return r.status, body
        return 200, (await (await tc.get(f"/gradio_api/call/{api}/{body['event_id']}")).read()).decode()

    async def main():
        dec = FakeDecoder()
        door = fd.FrontDoor(dec, window_s=0.05)
        async with TestClient(TestServer(fd.make_app(door))) as tc:
            a, b = np.arange(40), np.arange(100, 170)
            res = await asyncio.gather(call(tc, fd.API_STREAMING, a, [5, 40]), call(tc, fd.API_STREAMING, b, [5, 40]),
                                       call(tc, fd.API, a, [0.8]), call(tc, fd.API_STREAMING, a, [0, 40]))
        return dec, door, res, (a, b)
    dec, door, res, (a, b) = asyncio.run(main())
    assert [r[0] for r in res] == [200, 200, 200, 422] and "block_size" in res[3][1]["error"]
    for (status, text), t in zip(res[:3], (a, b, a)):
        wav = _wav_from_events(text)
        assert np.abs(wav - np.repeat(t.astype(np.float32) / 8192.0, 1920)).max() < 1e-4
    assert len(made) == 1 and made[0].args == (5, 40) and len(made[0].items) == 2      # both streaming requests in one scheduler
    assert dec.calls == [1]                                                             # the one-shot request on its own

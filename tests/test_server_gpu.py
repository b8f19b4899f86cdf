"""GPU: the transport with the real decoder behind it -- three concurrent WebSocket connections whose
tokens trickle in; each must receive the audio the offline batched scheduler produces for the same tokens (the NSF source noise
is drawn per call, so the comparison is the band-energy distance the other streaming tests use, plus exact lengths)."""
import asyncio

import numpy as np
import pytest
import torch

from parity_metrics import band_db

pytestmark = pytest.mark.gpu

aiohttp = pytest.importorskip("aiohttp")
from aiohttp.test_utils import TestClient, TestServer  # noqa: E402


async def _client(tc, server, tokens, burst, pause):
    ws = await tc.ws_connect("/api/chat")
    hello = await ws.receive_bytes()
    assert hello[0] == server.KIND_HANDSHAKE
    for k in range(0, len(tokens), burst):
        await ws.send_bytes(bytes([server.KIND_TOKENS]) + np.asarray(tokens[k:k + burst], "<i4").tobytes())
        await asyncio.sleep(pause)
    await ws.send_bytes(bytes([server.KIND_END]))
    pcm = []
    while True:
        msg = await asyncio.wait_for(ws.receive_bytes(), 120)
        if msg[0] == server.KIND_END:
            break
        pcm.append(np.frombuffer(msg[1:], "<i2"))
    await ws.close()
    return np.concatenate(pcm).astype(np.float32) / 32767.0


def test_server_streams_match_offline_scheduler():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from moss_speech_decoder_cosy_b200 import server, streaming
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    dec = AudioDecoder(seed=0)
    g = torch.Generator().manual_seed(7)
    P = 37
    voice = (torch.randint(0, 16384, (1, P), generator=g).int(),
             torch.clamp(torch.randn(1, 4 * P, 80, generator=g) * 2 - 5, min=-11.5129),
             torch.randn(1, 192, generator=g))
    toks = [torch.randint(0, 16384, (n,), generator=g).tolist() for n in (58, 41, 73)]

    offline = streaming.StreamScheduler(dec, 5, 40)
    for t in toks:
        offline.add(t, *voice)
    want = [w.float().cpu().reshape(-1).numpy() for w in offline.run()]

    async def main():
        sch = streaming.StreamScheduler(dec, 5, 40)
        state = server.ServerState(sch, voice, codec=server.Pcm16Codec)
        async with TestClient(TestServer(server.make_app(state))) as tc:
            got = await asyncio.gather(*[_client(tc, server, t, 5, 0.005) for t in toks])
            await asyncio.sleep(0.05)
        return state, sch, got
    state, sch, got = asyncio.run(main())
    assert state.max_batch > 1, "connections never shared a decode step"
    for k, (w, pcm) in enumerate(zip(want, got)):
        n = w.shape[0]
        assert pcm.shape[0] == -(-n // server.FRAME_SIZE) * server.FRAME_SIZE and not pcm[n:].any()
        assert np.isfinite(pcm).all() and np.abs(pcm[:n]).max() > 1e-3
        d = band_db(torch.from_numpy(np.ascontiguousarray(pcm[:n])).reshape(1, -1), torch.from_numpy(np.clip(w, -1, 1)).reshape(1, -1))
        print(f"connection {k}: {len(toks[k])} tokens, {n} samples, band distance vs offline scheduler {d:.2f} dB")
        assert d < 3.0
    assert sorted(sch._free) == [0, 1, 2]


def test_http_frontdoor_batches_requests_through_the_decoder():
    """The Gradio REST front door (http_frontdoor.py) with the real decoder: three concurrent client_streaming.sh-style requests
    share one decode_batch call; each gets 1920 samples per token, the band-energy distance to a direct decode_batch of the same
    inputs is at the NSF-noise floor.  The reference `.npz` carries the 24 kHz prompt waveform: its mel comes from feat_extractor."""
    import base64
    import io
    import json
    import re
    import wave
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from aiohttp import FormData
    from moss_speech_decoder_cosy_b200 import http_frontdoor as fd
    from moss_speech_decoder_cosy_b200.flow_inference import AudioDecoder
    from oracle import cases
    dec = AudioDecoder(seed=0)
    g = torch.Generator().manual_seed(9)
    P = 20
    prompt_wav = cases.prompt_waveform(4242, P * 4 * 480 + 700).reshape(-1).numpy()
    ref = io.BytesIO()
    np.savez(ref, prompt_token=torch.randint(0, 16384, (P,), generator=g).numpy(), embedding=torch.randn(192, generator=g).numpy(),
             prompt_wav=prompt_wav)
    toks = [torch.randint(0, 16384, (n,), generator=g).numpy() for n in (30, 52, 41)]
    door = fd.FrontDoor(dec, window_s=0.05)

    async def one(tc, t):
        form = FormData()
        b = io.BytesIO()
        np.save(b, t)
        form.add_field("files", b.getvalue(), filename="tokens.npy")
        form.add_field("files", ref.getvalue(), filename="ref.npz")
        paths = await (await tc.post("/gradio_api/upload", data=form)).json()
        r = await tc.post(f"/gradio_api/call/{fd.API}", json={"data": [{"path": paths[0]}, {"path": paths[1]}, 0.8]})
        ev = (await r.json())["event_id"]
        return (await (await tc.get(f"/gradio_api/call/{fd.API}/{ev}")).read()).decode()

    async def main():
        async with TestClient(TestServer(fd.make_app(door))) as tc:
            return await asyncio.gather(*[one(tc, t) for t in toks])
    texts = asyncio.run(main())
    assert sum(door.batches) == 3 and max(door.batches) > 1, door.batches
    # the same inputs straight through the public API
    voice = door.reference(next(p for p in [__import__("os").path.join(door.upload_dir, f) for f in __import__("os").listdir(door.upload_dir)]
                                if p.endswith("ref.npz")))
    direct = dec.decode_batch([torch.from_numpy(t).reshape(1, -1) for t in toks], [voice[0]] * 3, [voice[1]] * 3, voice[2].repeat(3, 1))
    dw, off = direct["wav"].float().cpu(), 0
    for t, text in zip(toks, texts):
        line = [l for l in text.splitlines() if re.match(r"^data: \[.+\]$", l)][0]
        html = json.loads(line[6:])[0]
        with wave.open(io.BytesIO(base64.b64decode(re.search(r"data:audio/wav;base64,([A-Za-z0-9+/=]+)", html).group(1)))) as w:
            pcm = np.frombuffer(w.readframes(w.getnframes()), "<i2").astype(np.float32) / 32767.0
        n = len(t) * 1920
        assert pcm.shape == (n,) and np.isfinite(pcm).all() and np.abs(pcm).max() > 1e-3
        d = band_db(torch.from_numpy(pcm).reshape(1, -1), dw[off:off + n].clamp(-1, 1).reshape(1, -1))
        print(f"front door, {len(t)} tokens: band distance vs direct decode_batch {d:.2f} dB")
        assert d < 3.0
        off += n

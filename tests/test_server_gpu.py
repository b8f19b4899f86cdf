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

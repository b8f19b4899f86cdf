"""Transport: the reference server.py wire protocol with the streaming scheduler behind it.
CPU only: the scheduler's decode step is replaced by a deterministic stand-in, everything else (incremental streams, chunk
schedule, cross-connection batching, framing, codec, handshake) is the shipped code."""
import asyncio

import numpy as np
import pytest
import torch

from moss_speech_decoder_cosy_b200 import server, streaming

aiohttp = pytest.importorskip("aiohttp")
from aiohttp.test_utils import TestClient, TestServer  # noqa: E402

SAMPLES_PER_TOKEN = 960            # 2 mel frames x 480 samples


class FakeScheduler(streaming.StreamScheduler):
    """StreamScheduler with the GPU work replaced: a chunk (start, end, actual, fin) emits one sample per generated
    sample position, valued token/8192, for the tokens the real decoder would emit audio for."""

    def __init__(self, block_size=5, max_token_len=40):
        super().__init__(dec=None, block_size=block_size, max_token_len=max_token_len, prefix_cache=False)
        self.batches = []

    def step(self):
        ready = {i: s.next_chunk() for i, s in enumerate(self.streams) if not s.done}
        ready = {i: c for i, c in ready.items() if c is not None}
        self.batches.append(sorted(ready))
        out = {}
        for i, c in ready.items():
            s = self.streams[i]
            start, end, actual, fin = c
            first = start + actual
            last = end if fin else end - streaming.PRE_LOOKAHEAD
            vals = np.asarray(s.tokens[first:last], np.float32) / 8192.0
            out[i] = torch.from_numpy(np.repeat(vals, SAMPLES_PER_TOKEN)).reshape(1, -1)
            s.out.append(out[i])
            s.advance(c)
        return out


def _voice(n_prompt=7):
    return (torch.zeros(1, n_prompt, dtype=torch.long), torch.zeros(1, 2 * n_prompt, 80), torch.zeros(1, 192))


def test_incremental_stream_follows_chunk_plan():
    for n, n_prompt, block, mtl in [(60, 38, 5, 40), (23, 0, 8, None), (33, 7, 5, 40), (4, 7, 5, 40), (200, 37, 5, 40)]:
        plan = streaming.chunk_plan(n, n_prompt, block, mtl)
        toks = list(range(n))
        # all tokens up front
        s = streaming.Stream(toks, torch.zeros(1, n_prompt), None, None, block, mtl)
        got = []
        while not s.done:
            c = s.next_chunk()
            got.append(c)
            s.advance(c)
        assert got == plan
        # tokens trickling in one at a time: same chunks, never an early one
        s = streaming.Stream([], torch.zeros(1, n_prompt), None, None, block, mtl, ended=False)
        got = []
        for t in toks:
            s.push([t])
            while (c := s.next_chunk()) is not None:
                got.append(c)
                s.advance(c)
        assert s.next_chunk() is None and not s.done
        s.finish()
        while (c := s.next_chunk()) is not None:
            got.append(c)
            s.advance(c)
        assert got == plan and s.done
        with pytest.raises(ValueError):
            s.push([1])


def test_scheduler_positions_are_recycled():
    sch = FakeScheduler()
    a = sch.add(list(range(30)), *_voice())
    b = sch.add(list(range(30)), *_voice())
    tag_a = sch.streams[a].tag
    sch.release(a)
    c = sch.add(list(range(12)), *_voice())
    assert c == a and sch.streams[c].tag != tag_a and not sch.streams[c].done
    wavs = sch.run()
    assert wavs[b].shape[-1] == 30 * SAMPLES_PER_TOKEN and wavs[c].shape[-1] == 12 * SAMPLES_PER_TOKEN


async def _client(tc, tokens, burst, pause):
    ws = await tc.ws_connect("/api/chat")
    hello = await ws.receive_bytes()
    assert hello[0] == server.KIND_HANDSHAKE and hello[1:] == b"pcm16"
    await ws.send_bytes(bytes([server.KIND_AUDIO]) + np.zeros(1920, "<i2").tobytes())     # microphone frame: accepted, dropped
    await ws.send_bytes(b"")                                                               # empty message: ignored
    await ws.send_bytes(bytes([9]) + b"zz")                                                # unknown kind: ignored
    for k in range(0, len(tokens), burst):
        await ws.send_bytes(bytes([server.KIND_TOKENS]) + np.asarray(tokens[k:k + burst], "<i4").tobytes())
        await asyncio.sleep(pause)
    await ws.send_bytes(bytes([server.KIND_END]))
    pcm = []
    while True:
        msg = await asyncio.wait_for(ws.receive_bytes(), 10)
        if msg[0] == server.KIND_END:
            break
        assert msg[0] == server.KIND_AUDIO and (len(msg) - 1) % (2 * server.FRAME_SIZE) == 0     # whole 1920-sample frames
        pcm.append(np.frombuffer(msg[1:], "<i2"))
    await ws.close()
    return np.concatenate(pcm)


def test_server_protocol_and_cross_connection_batching():
    async def main():
        sch = FakeScheduler()
        state = server.ServerState(sch, _voice(), codec=server.Pcm16Codec)
        async with TestClient(TestServer(server.make_app(state))) as tc:
            rng = np.random.default_rng(0)
            toks = [rng.integers(0, 8192, n).tolist() for n in (61, 33, 47)]
            got = await asyncio.gather(*[_client(tc, t, 7, 0.002) for t in toks])
            await asyncio.sleep(0.05)
        return sch, state, toks, got
    sch, state, toks, got = asyncio.run(main())
    for t, pcm in zip(toks, got):
        want = np.repeat(np.asarray(t, np.float32) / 8192.0, SAMPLES_PER_TOKEN)
        n = want.shape[0]
        padded = -(-n // server.FRAME_SIZE) * server.FRAME_SIZE
        assert pcm.shape[0] == padded                                   # tail frame zero padded
        assert np.abs(pcm[:n].astype(np.float32) / 32767.0 - want).max() < 1e-4
        assert not pcm[n:].any()
    assert any(len(b) > 1 for b in sch.batches), "connections never shared a scheduler step"
    assert state.max_batch > 1
    assert not state.conns and sorted(sch._free) == [0, 1, 2]       # every position handed back


def test_server_client_disconnect_cancels_stream():
    async def main():
        sch = FakeScheduler()
        state = server.ServerState(sch, _voice(), codec=server.Pcm16Codec)
        async with TestClient(TestServer(server.make_app(state))) as tc:
            ws = await tc.ws_connect("/api/chat")
            await ws.receive_bytes()
            await ws.send_bytes(bytes([server.KIND_TOKENS]) + np.arange(20, dtype="<i4").tobytes())
            await ws.close()
            await asyncio.sleep(0.1)
            assert not state.conns and sch.streams[0].done
            # the next connection takes the released position
            pcm = await _client(tc, list(range(12)), 12, 0.0)
            assert pcm.shape[0] == -(-12 * SAMPLES_PER_TOKEN // server.FRAME_SIZE) * server.FRAME_SIZE
            assert len(sch.streams) == 1
    asyncio.run(main())

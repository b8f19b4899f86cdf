"""Streaming half of the AudioDecoder mirror: `token2wav_streaming` (flow_inference.py:130-164), the chunk loop of
`stream_inference` (:166-243), and a batched scheduler that steps many concurrent streams through one flow solve and
one vocoder pass per chunk (BASELINE.json config 3; the reference serves one stream at a time).

Semantics kept from the reference: every chunk RECOMPUTES the flow over prompt + window tokens with block-causal
masks (`stream=True`), crops `token_offset * 4` frames, prepends the cached `mel_cache_len` mel frames, feeds the
cached source samples to HiFT (generator.py:579-580), cross-fades the first `source_cache_len` samples with the cached
speech tail under a Hamming window (flow_inference.py:35-43) and holds back the new tail.  The fade runs on the device
(the reference round-trips through the CPU and a float64 numpy window, flow_inference.py:37-43).
"""
import uuid as _uuid

import numpy as np
import torch

TOKEN_MEL_RATIO = 4
PRE_LOOKAHEAD = 3


def fade_in_out_mel(fade_in_mel, fade_out_mel, window):
    """flow_inference.py:35-43 for mel tensors (only reachable when a caller fills mel_overlap_dict)."""
    n = int(window.shape[0] / 2)
    w = torch.from_numpy(np.asarray(window, dtype=np.float32)).to(fade_in_mel.device)
    out = fade_in_mel.clone()
    out[..., :n] = fade_in_mel[..., :n] * w[:n] + fade_out_mel[..., -n:].to(out.device) * w[n:]
    return out


def _hift_chunks(dec, mels, caches, finalize, nsf_noise=None):
    """Shared tail of token2wav_streaming for a list of streams.

    mels[b]: time-major [T_b, 80] new mel frames (already cropped by token_offset); caches[b]: None or the dict
    {'mel' [1,80,8], 'source' [1,1,3840], 'speech' [1,3840]}.  Returns (speech list [1, n_b], new caches)."""
    dev = dec.device
    B = len(mels)
    full, lens, clen = [], [], []
    S = dec.source_cache_len
    cache_src = torch.zeros(B, S, device=dev)
    for b in range(B):
        m = mels[b]
        if caches[b] is not None:
            m = torch.cat([caches[b]['mel'][0].t().to(dev), m], 0)
            cs = caches[b]['source'].reshape(-1)
            cache_src[b, :cs.numel()] = cs
            clen.append(int(cs.numel()))
        else:
            clen.append(0)
        full.append(m)
        lens.append(int(m.shape[0]))
    mel_cat = torch.cat(full, 0)
    wav, src = dec.engine.hift(lens, mel_cat.reshape(-1), cache_source=cache_src, cache_len=clen,
                               nsf_noise=dec._noise_for(lens, nsf_noise),
                               seed=dec._next_seed(), time_major=True)
    out, new_caches, off = [], [], 0
    for b in range(B):
        n = lens[b] * 480
        speech, source = wav[off:off + n], src[off:off + n]
        off += n
        if caches[b] is not None:
            prev = caches[b]['speech'].reshape(-1)
            if int(prev.numel()) != S or n < S:
                # the reference's fade_in_out (flow_inference.py:35-43) fails with a shape error here (a previous chunk
                # shorter than mel_cache_len frames); a shorter fade would need differently aligned window halves
                raise ValueError(f"cross-fade needs {S} cached and {S} new samples, got {int(prev.numel())} / {n}")
            # speech[:S] = speech[:S] * w[:S] + prev * w[S:]   (Hamming window of 2 * source_cache_len)
            dec.engine.fade_in_out(speech, prev.contiguous(), dec._speech_window_dev)
        if not finalize:
            new_caches.append({'mel': full[b][-dec.mel_cache_len:].t().unsqueeze(0).contiguous(),
                               'source': source[-S:].clone().view(1, 1, -1), 'speech': speech[-S:].clone().view(1, -1)})
            out.append(speech[:-S].view(1, -1))
        else:
            new_caches.append(caches[b])
            out.append(speech.view(1, -1))
    return out, new_caches


@torch.inference_mode()
def token2wav_streaming(dec, token, prompt_token, prompt_feat, token_offset, uuid, embedding, finalize=False,
                        stream=False, speed=1.0):
    """flow_inference.py:130-164."""
    if embedding.shape[0] == 0:
        embedding = torch.zeros(1, 192)
    r = dec.flow_batch([token], [prompt_token], [prompt_feat], embedding, streaming=stream, finalize=finalize)
    mel = r["mel"][r["prompt_frames"][0]:]                       # feat_return, time-major
    mel = mel[token_offset * TOKEN_MEL_RATIO:]
    cache = dec.hift_cache_dict[uuid]
    if finalize and speed != 1.0:
        assert cache is None, 'speed change only support non-stream inference mode'
        m = torch.nn.functional.interpolate(mel.t().unsqueeze(0), size=int(mel.shape[0] / speed), mode='linear')
        mel = m[0].t().contiguous()
    out, caches = _hift_chunks(dec, [mel], [cache], finalize)
    if not finalize:
        dec.hift_cache_dict[uuid] = caches[0]
    return out[0]


def chunk_plan(n_tokens, n_prompt, block_size, max_token_len, pre_lookahead=PRE_LOOKAHEAD):
    """The chunk schedule of stream_inference (flow_inference.py:186-232) as a pure function:
    [(start, end, actual_token_offset, finalize)] over the token list.  Shared by the single-stream loop, the batched
    scheduler and the CPU tests."""
    plan = []
    token_offset = 0
    hop = block_size
    pad = int(np.ceil(n_prompt / hop) * hop - n_prompt)
    while True:
        this_hop = hop + pad if token_offset == 0 else hop
        if n_tokens - token_offset >= this_hop + pre_lookahead:
            end = token_offset + this_hop + pre_lookahead
            if max_token_len is not None:
                start = max(0, end - max_token_len)
                actual = token_offset - start if start > 0 else token_offset
            else:
                start, actual = 0, token_offset
            plan.append((start, end, actual, False))
            token_offset += this_hop
        else:
            break
    if max_token_len is not None:
        start = max(0, n_tokens - max_token_len)
        plan.append((start, n_tokens, token_offset - start, True))
    else:
        plan.append((0, n_tokens, token_offset, True))
    return plan


@torch.inference_mode()
def stream_inference(dec, token, prompt_token, prompt_feat, embedding, block_size=8, max_token_len=None):
    """flow_inference.py:166-243: chunked decode of one utterance; returns the concatenated waveform on the CPU."""
    this_uuid = str(_uuid.uuid1())
    toks = token.reshape(-1).cpu().tolist()
    dec.tts_speech_token_dict[this_uuid], dec.llm_end_dict[this_uuid] = toks, True
    dec.hift_cache_dict[this_uuid] = None
    speechs = []
    for start, end, actual, fin in chunk_plan(len(toks), prompt_token.shape[1], block_size, max_token_len):
        chunk = torch.tensor(toks[start:end]).unsqueeze(0)
        sp = token2wav_streaming(dec, chunk, prompt_token, prompt_feat, actual, this_uuid, embedding,
                                 finalize=fin, stream=not fin)
        speechs.append(sp.cpu())
    del dec.hift_cache_dict[this_uuid]
    del dec.tts_speech_token_dict[this_uuid]
    del dec.llm_end_dict[this_uuid]
    return torch.cat(speechs, dim=-1).cpu()


class Stream:
    """One concurrent session of the batched scheduler.  Tokens may be given up front (`ended=True`, the offline
    `stream_inference` case) or pushed while the stream runs (`push` / `finish`: the server case, where the language model
    produces them); the chunk schedule is that of stream_inference (flow_inference.py:186-232) evaluated lazily --
    `next_chunk()` yields exactly the entries of `chunk_plan` for the tokens seen so far."""
    _next_tag = 1

    def __init__(self, tokens, prompt_token, prompt_feat, embedding, block_size, max_token_len, ended=True):
        self.tokens = [int(t) for t in tokens]
        self.prompt_token, self.prompt_feat, self.embedding = prompt_token, prompt_feat, embedding
        self.block_size, self.max_token_len = block_size, max_token_len
        self.ended = bool(ended)
        self.token_offset, self.finalized = 0, False
        self.pad = int(np.ceil(prompt_token.shape[1] / block_size) * block_size - prompt_token.shape[1])
        self.step = 0
        self.cache = None
        self.out = []
        self.slot, self.recorded = None, False            # frozen-prefix cache slot of this stream
        self.tag = Stream._next_tag                        # identity of its frozen inputs (prompt, speaker)
        Stream._next_tag += 1

    def push(self, tokens):
        if self.ended:
            raise ValueError("stream already finished")
        self.tokens.extend(int(t) for t in tokens)

    def finish(self):
        self.ended = True

    def cancel(self):
        """Drop the stream (client went away): nothing more is decoded for it."""
        self.ended = self.finalized = True

    def next_chunk(self):
        """(start, end, actual_token_offset, finalize) of the chunk that can be decoded now, or None (waiting for tokens)."""
        if self.finalized:
            return None
        n = len(self.tokens)
        this_hop = self.block_size + self.pad if self.token_offset == 0 else self.block_size
        if n - self.token_offset >= this_hop + PRE_LOOKAHEAD:
            end = self.token_offset + this_hop + PRE_LOOKAHEAD
            if self.max_token_len is not None:
                start = max(0, end - self.max_token_len)
                actual = self.token_offset - start if start > 0 else self.token_offset
            else:
                start, actual = 0, self.token_offset
            return (start, end, actual, False)
        if not self.ended:
            return None
        if self.max_token_len is not None:
            start = max(0, n - self.max_token_len)
            return (start, n, self.token_offset - start, True)
        return (0, n, self.token_offset, True)

    def advance(self, chunk):
        """Bookkeeping after `chunk` (a value of next_chunk) has been decoded."""
        if chunk[3]:
            self.finalized = True
        else:
            self.token_offset += self.block_size + self.pad if self.token_offset == 0 else self.block_size
        self.step += 1

    @property
    def done(self):
        return self.finalized


class StreamScheduler:
    """Steps many streams together: every `step()` takes the next chunk of each unfinished stream, runs ONE batched
    flow solve (block-causal for non-final chunks, full attention for final ones, like the reference's per-call flags)
    and ONE batched vocoder pass, and returns {stream index: speech chunk [1, n]} on the device.

    prefix_cache: the reference recomputes the flow over prompt + window for every chunk
    (flow_inference.py:188-218).  The first whole attention chunks inside the prompt are identical in every chunk call of
    a stream, so their K | V and conv context are recorded by the stream's first chunk and reused afterwards
    (include/mossdec.h, mossdec_token2mel_prefix): exact, ~1/3 less work per chunk at a 37-token prompt + 40-token window."""

    def __init__(self, dec, block_size=5, max_token_len=40, prefix_cache=True):
        self.dec, self.block_size, self.max_token_len = dec, block_size, max_token_len
        self.streams = []
        self.prefix_cache = bool(prefix_cache)
        self._cap, self._prefix_frames = 0, 0
        self._free = []                                    # indices of released streams, reused by add()

    def add(self, tokens, prompt_token, prompt_feat, embedding, ended=True):
        """New stream; with ended=False more tokens follow through streams[i].push(...) / .finish()."""
        s = Stream(tokens, prompt_token, prompt_feat, embedding, self.block_size, self.max_token_len, ended)
        if self._free:                                     # a released position: its prefix-cache slot is re-recorded (new tag)
            i = self._free.pop()
            self.streams[i] = s
            return i
        self.streams.append(s)
        return len(self.streams) - 1

    def release(self, i):
        """Give position i (and its prefix-cache slot) back; a long-running server calls this when a connection ends."""
        self.streams[i].cancel()
        self.streams[i].out = []
        if i not in self._free:
            self._free.append(i)

    def _prefix_frames_for(self, s):
        """Frames of the frozen prefix a stream supports: whole estimator chunks (== whole encoder chunks) that lie inside the
        prompt mel AND whose tokens (+ the encoder's 3 look-ahead tokens) are prompt tokens."""
        dec = self.dec
        chunk = dec.estimator_chunk_frames
        if chunk <= 0 or chunk != dec.static_chunk_size * TOKEN_MEL_RATIO:
            return 0
        usable = min(int(s.prompt_feat.shape[1]), TOKEN_MEL_RATIO * (int(s.prompt_token.shape[1]) - PRE_LOOKAHEAD))
        return max(0, usable // chunk) * chunk

    def _ensure_cache(self):
        """One cache geometry per scheduler: the smallest prefix any eligible stream supports (>= one chunk)."""
        if not self.prefix_cache:
            return
        frames = [f for f in (self._prefix_frames_for(s) for s in self.streams) if f > 0]
        if not frames:
            return
        want = min(frames)
        n = len(self.streams)
        if want != self._prefix_frames or n > self._cap:
            self._cap = max(n, 2 * self._cap, 4)
            self._prefix_frames = want
            self.dec.engine.prefix_reserve(self._cap, want, self.dec.n_timesteps)
            for i, s in enumerate(self.streams):               # (re)allocation drops the contents
                s.slot, s.recorded = None, False
        for i, s in enumerate(self.streams):
            if s.slot is None and self._prefix_frames_for(s) >= self._prefix_frames:
                s.slot = i

    @torch.inference_mode()
    def step(self):
        dec = self.dec
        emitted = {}
        self._ensure_cache()
        ready = {i: s.next_chunk() for i, s in enumerate(self.streams) if not s.done}
        ready = {i: c for i, c in ready.items() if c is not None}
        for fin in (False, True):
            live = [i for i, c in ready.items() if c[3] == fin]
            # non-final chunks: streams with a recorded prefix reuse it (mode 2), eligible new ones record it (mode 1)
            groups = {0: [], 1: [], 2: []}
            for i in live:
                s = self.streams[i]
                mode = 0 if (fin or s.slot is None) else (2 if s.recorded else 1)
                groups[mode].append(i)
            for mode, ids in groups.items():
                if not ids:
                    continue
                toks, offs = [], []
                for i in ids:
                    s = self.streams[i]
                    start, end, actual, _ = ready[i]
                    toks.append(torch.tensor(s.tokens[start:end]).unsqueeze(0))
                    offs.append(actual)
                emb = torch.cat([self.streams[i].embedding.reshape(1, -1) for i in ids], 0)
                prefix = None
                if mode:
                    prefix = {"mode": mode, "slots": [self.streams[i].slot for i in ids],
                              "tags": [self.streams[i].tag for i in ids]}
                r = dec.flow_batch(toks, [self.streams[i].prompt_token for i in ids],
                                   [self.streams[i].prompt_feat for i in ids], emb, streaming=not fin, finalize=fin,
                                   prefix=prefix)
                mels, off = [], 0
                for k in range(len(ids)):
                    T, F_ = r["frames"][k], r["prompt_frames"][k]
                    mels.append(r["mel"][off + F_ + offs[k] * TOKEN_MEL_RATIO: off + T])
                    off += T
                out, caches = _hift_chunks(dec, mels, [self.streams[i].cache for i in ids], fin)
                for k, i in enumerate(ids):
                    s = self.streams[i]
                    s.cache = caches[k]
                    s.advance(ready[i])
                    s.out.append(out[k])
                    emitted[i] = out[k]
                    if mode == 1:
                        s.recorded = True
        return emitted

    def run(self):
        while any(not s.done for s in self.streams):
            self.step()
        return [torch.cat(s.out, dim=-1) for s in self.streams]

"""Multi-GPU partitioning of independent utterances.

The reference scales by rank-striding a list of utterances with no communication
(benchmark_moss_decoder.py:175-189: `data[rank::world_size]`, one process per GPU).  Same model here: every rank owns a
full copy of the weights and decodes its share; there is no collective on the data path.  Beyond plain striding this
module balances mixed-length sets with the path's cost model and buckets each share into batches of similar length.
"""
from typing import List, Sequence

EST_FLOP_PER_FRAME = 132.162e6     # estimator GEMM-class FLOPs per frame, branch and Euler step
ATTN_FLOP_PER_FRAME_T = 0.114688e6  # attention FLOPs per frame, branch and step, times T
HIFT_FLOP_PER_FRAME = 612.296e6
ENC_FLOP_PER_FRAME = 56e6


def utterance_cost(n_tokens: int, n_prompt: int = 0, n_steps: int = 10) -> float:
    """Algorithmic FLOPs of token2wav for one utterance."""
    T = 4 * (n_tokens + n_prompt)
    return n_steps * 2 * T * (EST_FLOP_PER_FRAME + ATTN_FLOP_PER_FRAME_T * T) + ENC_FLOP_PER_FRAME * T + \
        HIFT_FLOP_PER_FRAME * 4 * n_tokens


def stride_partition(n_items: int, rank: int, world: int) -> List[int]:
    """The reference's split: indices rank, rank + world, ... (benchmark_moss_decoder.py:189)."""
    return list(range(rank, n_items, world))


def balanced_partition(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time-first: deterministic, identical on every rank (no communication needed)."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    shares: List[List[int]] = [[] for _ in range(world)]
    load = [0.0] * world
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        shares[r].append(i)
        load[r] += costs[i]
    return shares


def length_buckets(indices: Sequence[int], lengths: Sequence[int], max_frames: int, max_batch: int = 64) -> List[List[int]]:
    """Group a rank's utterances into batches of similar length, bounded by total frames and batch size."""
    order = sorted(indices, key=lambda i: (lengths[i], i))
    out, cur, frames = [], [], 0
    for i in order:
        f = 4 * lengths[i]
        if cur and (frames + f > max_frames or len(cur) >= max_batch):
            out.append(cur)
            cur, frames = [], 0
        cur.append(i)
        frames += f
    if cur:
        out.append(cur)
    return out

"""How the path is partitioned over ranks / GPUs (host-side arithmetic only).

Two partitions, both from BASELINE.json:
  * site blocks (configs[3]): rank g owns a contiguous block of the words of EVERY row — the
    reference's own OpenMP decomposition (/root/reference/src/TreeEvaluation.c:95-97, :114: thread t
    owns words [t*slice, (t+1)*slice), the last one also takes the remainder).  Every rank runs
    the same programs; partial tree lengths are summed (one all-reduce of K int64).
  * replicates (configs[4]): independent SA chains, chain r has seed s + r; ranks take
    contiguous ranges of chains and never talk on the data path.
  * candidates (third row): every rank holds the WHOLE alignment and the same current
    tree and scores a contiguous range of the K candidates of a batch (an NNI neighbourhood, the
    speculative proposals of a chain); the K lengths are gathered by the host layer, and the accepted
    candidate is committed by every rank itself (its dirty rows are recomputed locally: cheaper than
    shipping them).  No collective on the device side.
Used by bench.py, by the multi-rank tests, and mirrored in C++ by the host layer (-D).
"""
from __future__ import annotations

from typing import Tuple


def word_block(nwords: int, rank: int, world: int, align: int = 2) -> Tuple[int, int]:
    """(word_begin, word_count) of `rank`: contiguous, covering [0, nwords) exactly once, block
    boundaries on multiples of `align` words (2 words = one 16-byte device column) so that no
    column is split between ranks; the last rank takes the remainder, as in getplen's slices."""
    if not (0 <= rank < world) or nwords < 1:
        raise ValueError("bad rank/world/nwords")
    units = (nwords + align - 1) // align
    if units < world:
        raise ValueError(f"{world} ranks are too many for {nwords} words: a rank would own nothing")

    def start(r: int) -> int:
        return min(nwords, (units * r // world) * align)

    b, e = start(rank), (nwords if rank == world - 1 else start(rank + 1))
    return b, e - b


def chain_range(replicates: int, rank: int, world: int) -> Tuple[int, int]:
    """(first_chain, count) of `rank`: contiguous ranges, sizes differ by at most one."""
    if not (0 <= rank < world) or replicates < 0:
        raise ValueError("bad rank/world/replicates")
    b = replicates * rank // world
    e = replicates * (rank + 1) // world
    return b, e - b


def replicate_shard(seed: int, replicates: int, rank: int, world: int) -> Tuple[int, int]:
    """(seed of the rank's first chain, number of chains): chain r always runs with seed + r."""
    first, count = chain_range(replicates, rank, world)
    return seed + first, count


def candidate_range(candidates: int, rank: int, world: int) -> Tuple[int, int]:
    """(first_candidate, count) of `rank` for a batch of `candidates`: contiguous, sizes differ by at most
    one, every candidate scored exactly once."""
    return chain_range(candidates, rank, world)

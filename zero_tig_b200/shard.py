"""Frame-range sharding of clips across the GPUs of one box.

The temporal state is a true recurrence (frame t's H3/s3 feed frame t+1, model/model.py:324-339), so the
exact shard unit is a *sequence* (boundaries as utils.sequential_judgment, utils/utils.py:145-160:
directory change or non-consecutive index).  Whole sequences are dealt to ranks first; a rank that would
otherwise idle takes a contiguous frame range of a long sequence, preceded by ONE halo frame that is
processed as ``is_new_seq=True`` only to seed ``last_H3/last_s3`` and whose output is discarded.  A
sharded run therefore equals the reference run independently on each (halo + range) sub-clip.
No data-path collective is involved.
"""
import os
from dataclasses import dataclass
from typing import List, Sequence, Tuple


@dataclass(frozen=True)
class Work:
    """Frames [start, stop) of sequence `seq` for one rank; `halo` = index of the seed frame or -1."""
    seq: int
    start: int
    stop: int
    halo: int


def sequence_lengths(paths: Sequence[str]) -> List[int]:
    """Split an ordered list of frame paths into sequences with the reference's rule."""
    out, prev = [], None
    for p in paths:
        d, name = os.path.split(p)
        idx = int(os.path.splitext(name)[0])
        if prev is None or d != prev[0] or idx != prev[1] + 1:
            out.append(0)
        out[-1] += 1
        prev = (d, idx)
    return out


def plan(seq_lengths: Sequence[int], world: int) -> List[List[Work]]:
    """Balanced plan: cut the concatenated frame axis into `world` equal ranges, never across a sequence
    boundary without starting a fresh sequence, adding a halo frame where a range starts mid-sequence."""
    total = sum(seq_lengths)
    plans: List[List[Work]] = [[] for _ in range(world)]
    if total == 0:
        return plans
    bounds = [round(total * r / world) for r in range(world + 1)]
    base = 0
    for s, n in enumerate(seq_lengths):
        for r in range(world):
            lo, hi = max(bounds[r], base), min(bounds[r + 1], base + n)
            if lo < hi:
                start = lo - base
                plans[r].append(Work(s, start, hi - base, start - 1 if start > 0 else -1))
        base += n
    return plans


def run(model, frames_of_seq, works: Sequence[Work], emit) -> int:
    """Drive `model` (a Finetunemodel-like callable with `.is_new_seq`) over one rank's plan.
    frames_of_seq(seq, idx) returns the input tensor; emit(seq, idx, outputs) receives the kept results.
    Returns the number of model calls (kept frames + halos)."""
    calls = 0
    for w in works:
        if w.halo >= 0:
            model.is_new_seq = True
            model(frames_of_seq(w.seq, w.halo))          # seed only, output discarded
            calls += 1
        for i in range(w.start, w.stop):
            model.is_new_seq = (i == w.start and w.halo < 0)
            emit(w.seq, i, model(frames_of_seq(w.seq, i)))
            calls += 1
    return calls


def coverage(plans: Sequence[Sequence[Work]]) -> List[Tuple[int, int]]:
    return sorted((w.seq, i) for p in plans for w in p for i in range(w.start, w.stop))

"""CPU: frame-range sharding logic, including a world_size-2 gloo run of the N>1 host path."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from zero_tig_b200 import shard


def test_sequence_lengths_follow_reference_rule():
    paths = ["a/00001.png", "a/00002.png", "a/00004.png", "b/00005.png", "b/00006.png"]
    assert shard.sequence_lengths(paths) == [2, 1, 2]


def test_plan_covers_every_frame_once_with_halos():
    for lens, world in (([64], 8), ([5, 7, 64, 3], 4), ([3], 8), ([10, 10], 2), ([], 2)):
        plans = shard.plan(lens, world)
        assert shard.coverage(plans) == [(s, i) for s, n in enumerate(lens) for i in range(n)]
        for p in plans:
            for w in p:
                assert w.halo == (w.start - 1 if w.start > 0 else -1)
        sizes = [sum(w.stop - w.start for w in p) for p in plans]
        assert max(sizes) - min(sizes) <= 1 or sum(lens) < world


class _FakeModel:
    def __init__(self):
        self.is_new_seq, self.log = True, []

    def __call__(self, x):
        self.log.append((int(x), self.is_new_seq))
        return int(x)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lens = [9, 4]
    works = shard.plan(lens, world)[rank]
    m, kept = _FakeModel(), []
    calls = shard.run(m, lambda s, i: 100 * s + i, works, lambda s, i, o: kept.append(o))
    # fps-style bookkeeping is the only cross-rank traffic: a MAX over ranks of the per-rank work
    t = torch.tensor([float(calls)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gathered = [None] * world
    dist.all_gather_object(gathered, (kept, m.log))
    if rank == 0:
        q.put((gathered, t.item()))
    dist.destroy_process_group()


def test_two_rank_gloo_run():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, max_calls = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    kept = sorted(k for g in gathered for k in g[0])
    assert kept == list(range(9)) + [100 + i for i in range(4)]
    log1 = gathered[1][1]
    # rank 1 starts mid-sequence 0: one halo frame processed as a new sequence, then steady frames
    assert log1[0] == (5, True) and log1[1] == (6, False)
    assert (100, True) in log1            # sequence 1 starts fresh
    assert max_calls == 8.0

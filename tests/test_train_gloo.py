"""CPU: the data-parallel host path of the zero-shot training step on a world_size-2 gloo
group: one sum all-reduce of the flat gradient buffer, 1/world folded into the clip scale, identical
buffers on every rank afterwards.  The clip+Adam kernel itself is CUDA-only (tests/test_gpu_train.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from zero_tig_b200.optim import ClipAdam
    g = torch.Generator().manual_seed(0)
    params = [torch.nn.Parameter(torch.randn(s, generator=g)) for s in ((4, 3, 3, 3), (4,), (2, 4, 1, 1))]
    before = [p.detach().clone() for p in params]
    opt = ClipAdam(params)
    assert all(torch.equal(p, b) for p, b in zip(params, before))          # flattening keeps the values
    assert opt.flat.numel() == sum(p.numel() for p in params) == 4 * 27 + 4 + 8
    opt.zero_grad()
    params[0].grad.add_(float(rank + 1))                 # in place into the flat views
    params[1].grad = torch.full((4,), 10.0 * (rank + 1))   # replaced tensor: gathered before the collective
    params[2].grad = None                                # missing gradient counts as zero
    opt.gather_grads()
    scale = opt.reduce_grads()
    out = (scale, opt.grad.clone(), opt.world_size())
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    if rank == 0:
        q.put(gathered)
    dist.destroy_process_group()


def test_flat_gradient_allreduce_two_ranks():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (s0, g0, w0), (s1, g1, w1) = gathered
    assert w0 == w1 == 2 and s0 == s1 == 0.5
    assert torch.equal(g0, g1)
    assert torch.all(g0[:108] == 3.0) and torch.all(g0[108:112] == 30.0) and torch.all(g0[112:] == 0.0)


def test_single_process_has_no_collective():
    from zero_tig_b200.optim import ClipAdam
    p = torch.nn.Parameter(torch.ones(5))
    opt = ClipAdam([p])
    p.grad.fill_(2.0)
    assert opt.reduce_grads() == 1.0 and torch.all(opt.grad == 2.0)

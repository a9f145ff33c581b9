"""GPU: the persistent RAFT update kernel (csrc/raft_update.cu: all iterations of raft.py:111-120 in one cooperative
launch, TMA-fed tcgen05 phases + CUDA-core lookup / flow head) against the round-1 launch chain (conv_tc + lookup +
flow-head kernels, itself gated against the oracle and the reference goldens in test_gpu_e2e.py).  Both paths feed
the tensor cores the same bf16 hi/lo operands in the same K order, so they agree to fp32 rounding; the end-to-end
gates (EPE <= 0.01 px vs the reference) run on the fused path by default."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(weights):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.RAFT.raft import RAFT
    raft = RAFT(synth.make_args()).cuda().eval()
    raft.load_state_dict({k[5:]: v for k, v in weights.items() if k.startswith("raft.")})
    return raft.engine()


def _inputs(hp, wp, seed):
    g = torch.Generator().manual_seed(seed)
    base = torch.rand(1, hp + 16, wp + 16, 3, generator=g)
    base = torch.nn.functional.avg_pool2d(base.permute(0, 3, 1, 2), 5, 1, 2).permute(0, 2, 3, 1)   # smooth texture
    a = base[:, 4:4 + hp, 6:6 + wp]
    b = base[:, 6:6 + hp, 3:3 + wp]                       # shifted by (2, -3) px: a real flow to find
    x = torch.cat([a, b], 0) * 2 - 1
    return torch.nn.functional.pad(x, (0, 1)).contiguous().cuda()          # [2][hp][wp][4]


# grids: 16x20 (one row of tiles, partial last column tile), 23x40 (configs[2]: partial rows), 45x80 (1080p), 18x17 (odd)
@pytest.mark.parametrize("hp,wp,iters,ctas", [(128, 160, 12, 0), (184, 320, 12, 0), (360, 640, 12, 0), (144, 136, 3, 0),
                                              (184, 320, 4, 7), (128, 160, 1, 0), (360, 640, 2, 33)])
def test_fused_update_matches_launch_chain(engine, hp, wp, iters, ctas):
    x2 = _inputs(hp, wp, hp + iters)
    engine.UPD_FUSED, engine.FUSED_MAX_CTAS = 0, 0
    ref_up = engine.forward_packed(x2, hp, wp, iters).clone()
    ref_flow = engine.last["flow4"][:, :2].clone()
    ref_h = engine.ws.get("hx", (engine.last["N"], 384))[:, :128].clone()
    engine.UPD_FUSED, engine.FUSED_MAX_CTAS = 1, ctas
    try:
        got_up = engine.forward_packed(x2, hp, wp, iters).clone()
        got_flow = engine.last["flow4"][:, :2].clone()
        got_h = engine.last["h"].clone()
    finally:
        engine.UPD_FUSED, engine.FUSED_MAX_CTAS = 1, 0
    torch.cuda.synchronize()
    assert torch.isfinite(got_up).all()
    assert ref_flow.abs().max().item() > 0.05             # the pair really moves (random weights: a small flow)
    assert (got_flow - ref_flow).abs().max().item() < 2e-4, (got_flow - ref_flow).abs().max().item()
    assert (got_up - ref_up).abs().max().item() < 2e-3
    assert (got_h - ref_h).abs().max().item() < 2e-4


def test_fused_update_is_deterministic(engine):
    hp, wp = 184, 320
    x2 = _inputs(hp, wp, 5)
    engine.UPD_FUSED = 1
    a = engine.forward_packed(x2, hp, wp, 12).clone()
    b = engine.forward_packed(x2, hp, wp, 12).clone()
    assert torch.equal(a, b)


def test_split_bf16_hilo():
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(1)
    x = (torch.randn(37, 24, generator=g) * 3).cuda()
    hi = torch.empty(37, 16, dtype=torch.bfloat16, device="cuda")
    lo = torch.empty_like(hi)
    ops.split_bf16_hilo(x, 24, hi, lo, 16, 16, 37)
    want_hi = x[:, :16].to(torch.bfloat16)
    assert torch.equal(hi, want_hi)
    assert torch.equal(lo, (x[:, :16] - want_hi.float()).to(torch.bfloat16))
    assert ((hi.float() + lo.float()) - x[:, :16]).abs().max().item() < 3e-5 * 3 * 4


def test_bf16_corr_storage_arm(engine):
    """configs[4] second arm: the whole RAFT pass with the pyramid STORED in bf16 (ZT_CORR_BF16) against the fp32-storage
    pass on the same pair; the flow difference is reported (profiles/) - it is an option, not the default path."""
    hp, wp = 360, 640
    x2 = _inputs(hp, wp, 77)
    engine.CORR_BF16 = 0
    a = engine.forward_packed(x2, hp, wp, 12).clone()
    engine.CORR_BF16 = 1
    try:
        b = engine.forward_packed(x2, hp, wp, 12).clone()
        assert engine.last["levels"][0].dtype == torch.bfloat16
    finally:
        engine.CORR_BF16 = 0
    d = torch.sqrt(((a - b) ** 2).sum(1)).mean().item()
    print("bf16-storage pyramid vs fp32-storage: mean EPE %.5f px, max |d| %.5f px" % (d, (a - b).abs().max().item()))
    assert torch.isfinite(b).all() and d < 0.1

"""GPU: every C-ABI kernel against the CPU oracle (oracle/zt_oracle.py) and the reference
golden vectors (tests/golden/ops.npz), on seeded inputs.  Integer work must be bit exact;
fp32 tolerances are written beside each check."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "ops.npz"))


def nhwc(x, cs=None):
    """NCHW cpu -> NHWC cuda [B*H*W][cs]"""
    B, C, H, W = x.shape
    cs = cs or (C + 3) // 4 * 4
    o = torch.zeros(B, H, W, cs)
    o[..., :C] = x.permute(0, 2, 3, 1)
    return o.reshape(B * H * W, cs).cuda().contiguous()


def nchw(y, B, H, W, C):
    return y.reshape(B, H, W, -1)[..., :C].permute(0, 3, 1, 2).cpu()


CONV_CASES = [
    # cin, cout, kh, kw, stride, pad, H, W, B
    (3, 64, 7, 7, 2, (3, 3), 40, 48, 2),
    (64, 96, 3, 3, 2, (1, 1), 20, 24, 2),
    (64, 96, 1, 1, 2, (0, 0), 20, 24, 1),
    (384, 128, 1, 5, 1, (0, 2), 9, 13, 1),
    (384, 128, 5, 1, 1, (2, 0), 9, 13, 1),
    (324, 256, 1, 1, 1, (0, 0), 9, 13, 1),
    (256, 126, 3, 3, 1, (1, 1), 9, 13, 1),
    (256, 2, 3, 3, 1, (1, 1), 9, 13, 1),
    (2, 128, 7, 7, 1, (3, 3), 9, 13, 1),
    (48, 48, 3, 3, 1, (1, 1), 70, 90, 1),
    (64, 64, 3, 3, 1, (1, 1), 130, 150, 1),
    (256, 576, 1, 1, 1, (0, 0), 9, 13, 1),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv2d_matches_torch(case, dev):
    from zero_tig_b200 import ops
    cin, cout, kh, kw, s, pad, H, W, B = case
    g = torch.Generator().manual_seed(hash(case) % 1000)
    x = torch.randn(B, cin, H, W, generator=g)
    w = torch.randn(cout, cin, kh, kw, generator=g) / (cin * kh * kw) ** 0.5
    b = torch.randn(cout, generator=g)
    ref = F.leaky_relu(F.conv2d(x.double(), w.double(), b.double(), stride=s, padding=pad), 0.2)
    wt = ops.ConvW(w.cuda(), b.cuda())
    Ho, Wo = ref.shape[-2:]
    out_cs = (cout + 3) // 4 * 4
    out = torch.zeros(B * Ho * Wo, out_cs, device=dev)
    xin = nhwc(x)
    ops.conv2d(wt, xin, xin.shape[1], wt.cin, B, H, W, out, out_cs, stride=s, pad=pad, act="lrelu")
    torch.cuda.synchronize()
    got = nchw(out, B, Ho, Wo, cout)
    # fp32 accumulation over K <= 1920 terms of O(1) products: abs 2e-5 is ~30 ulp of the result scale
    assert (got.double() - ref).abs().max().item() < 2e-5


def test_conv2d_two_sources_and_gru_epilogues(dev):
    """virtual cat + fused GRU gates == SepConvGRU horizontal half (update.py:45-52)."""
    from zero_tig_b200 import ops
    import zt_oracle as O
    g = torch.Generator().manual_seed(5)
    H, W = 10, 14
    h = torch.tanh(torch.randn(1, 128, H, W, generator=g))
    x = torch.randn(1, 256, H, W, generator=g)
    wd = {}
    for n in "zrq":
        for s in "12":
            wd[f"g.conv{n}{s}.weight"] = torch.randn(128, 384, *((1, 5) if s == "1" else (5, 1)), generator=g) * 0.03
            wd[f"g.conv{n}{s}.bias"] = torch.randn(128, generator=g) * 0.1
    ref = O.sep_conv_gru(h, x, wd, "g")
    hx = nhwc(torch.cat([h, x], 1))
    z = torch.empty(H * W, 128, device=dev)
    rh = torch.empty(H * W, 128, device=dev)
    for s, pad in (("1", (0, 2)), ("2", (2, 0))):
        wzr = ops.ConvW(torch.cat([wd[f"g.convz{s}.weight"], wd[f"g.convr{s}.weight"]]).cuda(),
                        torch.cat([wd[f"g.convz{s}.bias"], wd[f"g.convr{s}.bias"]]).cuda())
        wq = ops.ConvW(wd[f"g.convq{s}.weight"].cuda(), wd[f"g.convq{s}.bias"].cuda())
        ops.conv2d(wzr, hx, 384, 384, 1, H, W, rh, 128, pad=pad, epi="gru_zr", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128)
        ops.conv2d(wq, rh, 128, 128, 1, H, W, hx, 384, in1=hx, in1_cs=384, c1=256, in1_off=128, pad=pad,
                   epi="gru_q", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128)
    got = nchw(hx, 1, H, W, 128)
    assert (got - ref).abs().max().item() < 2e-5


def test_conv2d_residual_epilogues(dev):
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(6)
    x = torch.randn(1, 64, 20, 24, generator=g)
    w = torch.randn(64, 64, 3, 3, generator=g) * 0.05
    b = torch.randn(64, generator=g) * 0.1
    bn = [torch.rand(64, generator=g) + 0.5, torch.randn(64, generator=g) * 0.1,
          torch.randn(64, generator=g) * 0.1, torch.rand(64, generator=g) + 0.5]
    y = F.batch_norm(F.conv2d(x, w, b, padding=1), bn[2], bn[3], bn[0], bn[1], False, 0.0, 1e-5)
    wt = ops.ConvW(w.cuda(), b.cuda(), bn=[t.cuda() for t in bn])
    xin = nhwc(x)
    out = torch.empty_like(xin)
    ops.conv2d(wt, xin, 64, 64, 1, 20, 24, out, 64, pad=(1, 1), act="relu", epi="add", aux0=xin, aux0_cs=64)
    assert (nchw(out, 1, 20, 24, 64) - (x + F.relu(y))).abs().max().item() < 2e-5
    ops.conv2d(wt, xin, 64, 64, 1, 20, 24, out, 64, pad=(1, 1), act="relu", epi="add_relu", aux0=xin, aux0_cs=64)
    assert (nchw(out, 1, 20, 24, 64) - F.relu(x + F.relu(y))).abs().max().item() < 2e-5


@pytest.mark.parametrize("C", [64, 96, 128])
def test_instance_norm(C, dev):
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(C)
    x = torch.randn(2, C, 23, 31, generator=g) * 3 + 5   # large mean: exercises cancellation
    res = torch.randn(2, C, 23, 31, generator=g)
    ref = F.relu(res + F.relu(F.instance_norm(x.double())))
    xin, rin = nhwc(x), nhwc(res)
    out = torch.empty_like(xin)
    stats = torch.empty(2 * C * 2, dtype=torch.float64, device=dev)
    ops.instnorm(xin, C, out, C, stats, 2, 23 * 31, C, relu_inner=True, relu_outer=True, res=rin, res_cs=C)
    assert (nchw(out, 2, 23, 31, C).double() - ref).abs().max().item() < 5e-6


def _gpu_pyramid(fm1, fm2, dev):
    """corr volume + pooling through the same calls RaftEngine makes."""
    from zero_tig_b200 import ops
    _, D, h, w = fm1.shape
    N = h * w
    N4 = (N + 3) // 4 * 4
    f1, f2 = nhwc(fm1), nhwc(fm2)
    f2t = torch.empty(D, N4, device=dev)
    ops.transpose(f2, N, D, D, f2t, N4)
    hs, ws = [h], [w]
    for _ in range(3):
        hs.append(hs[-1] // 2)
        ws.append(ws[-1] // 2)
    lv = [torch.empty(N, hs[i] * ws[i], device=dev) for i in range(4)]
    cw = ops.ConvW.__new__(ops.ConvW)
    cw.w, cw.b, cw.kh, cw.kw, cw.cin, cw.cout, cw.w_cs = f2t, None, 1, 1, D, N, N4
    ops.conv2d(cw, f1, D, D, 1, h, w, lv[0], N, scale=0.0625)
    for i in range(3):
        if hs[i + 1] > 0 and ws[i + 1] > 0:
            ops.avgpool2(lv[i], lv[i + 1], N, hs[i], ws[i])
    return lv, hs, ws


def test_corr_volume_pyramid_vs_reference_golden(gold, dev):
    lv, hs, ws = _gpu_pyramid(torch.from_numpy(gold["corr_f1"]), torch.from_numpy(gold["corr_f2"]), dev)
    for i in range(4):
        ref = gold[f"corr_pyr{i}"].reshape(90, -1)
        # K=256 fp32 dot products of N(0,1) data (|corr| ~ 1): 2e-5 abs
        np.testing.assert_allclose(lv[i].cpu().numpy(), ref, rtol=0, atol=2e-5)


def _fused_pyramid(fm1, fm2, dev):
    """zt_corr_pyramid_f32: volume + three pooled levels from one tcgen05 accumulator."""
    from zero_tig_b200 import ops
    _, D, h, w = fm1.shape
    N = h * w
    fmap = torch.cat([nhwc(fm1), nhwc(fm2)], 0).contiguous()
    hs, ws = [h], [w]
    for _ in range(3):
        hs.append(hs[-1] // 2)
        ws.append(ws[-1] // 2)
    lv = [torch.full((N, max(hs[i] * ws[i], 1)), float("nan"), device=dev) for i in range(4)]
    scratch = torch.empty(ops.corr_pyramid_scratch_bytes(h, w) // 4, device=dev)
    ops.corr_pyramid(fmap, D, h, w, lv, scratch)
    return lv, hs, ws


def test_corr_pyramid_fused_vs_reference_golden(gold, dev):
    lv, hs, ws = _fused_pyramid(torch.from_numpy(gold["corr_f1"]), torch.from_numpy(gold["corr_f2"]), dev)
    for i in range(4):
        ref = gold[f"corr_pyr{i}"].reshape(90, -1)
        if ref.shape[1]:
            # bf16 hi/lo split (3 of 4 partial products), fp32 accumulation: 2e-5 abs on |corr| ~ 1
            np.testing.assert_allclose(lv[i].cpu().numpy(), ref, rtol=0, atol=2e-5)


@pytest.mark.parametrize("h,w", [(8, 8), (16, 16), (17, 23), (13, 9), (45, 80), (34, 61)])
def test_corr_pyramid_fused_vs_oracle(dev, h, w):
    import zt_oracle as O
    g = torch.Generator().manual_seed(h * 100 + w)
    f1 = torch.randn(1, 256, h, w, generator=g)
    f2 = torch.randn(1, 256, h, w, generator=g)
    pyr = O.corr_pyramid(f1, f2)
    lv, hs, ws = _fused_pyramid(f1, f2, dev)
    for i in range(4):
        ref = pyr[i].reshape(h * w, -1)
        if ref.shape[1]:
            got = lv[i].cpu()
            assert torch.isfinite(got).all(), f"level {i}: unwritten entries"
            # bf16 hi/lo split drops the lo*lo partial product (2^-16 relative per term); |corr| reaches ~5 here
            assert (got - ref).abs().max().item() < 5e-5, (i, (got - ref).abs().max().item())


@pytest.mark.parametrize("h,w", [(16, 16), (17, 23), (45, 80)])
def test_corr_pyramid_bf16_storage_is_the_rounded_fp32_pyramid(dev, h, w):
    """zt_corr_pyramid_bf16 (configs[4] "bf16 corr"): same GEMM, levels rounded to bf16 on the way out - every entry
    is within one bf16 rounding (2^-9 relative) of the fp32-storage kernel's, and the bf16 lookup returns exactly the
    fp32 lookup of those rounded planes."""
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(h + w)
    f1 = torch.randn(1, 256, h, w, generator=g)
    f2 = torch.randn(1, 256, h, w, generator=g)
    lv, hs, ws = _fused_pyramid(f1, f2, dev)
    N = h * w
    fmap = torch.cat([nhwc(f1), nhwc(f2)], 0).contiguous()
    lb = [torch.full((N, max(hs[i] * ws[i], 1)), float("nan"), device=dev).to(torch.bfloat16) for i in range(4)]
    scratch = torch.empty(ops.corr_pyramid_scratch_bytes(h, w) // 4, device=dev)
    ops.corr_pyramid(fmap, 256, h, w, lb, scratch)
    for i in range(4):
        if hs[i] * ws[i]:
            assert torch.isfinite(lb[i].float()).all()
            assert torch.equal(lb[i], lv[i].to(torch.bfloat16)), i
    if min(hs[3], ws[3]) >= 2:
        coords = (torch.rand(N, 2, generator=g) * torch.tensor([w + 6.0, h + 6.0]) - 3).cuda()
        a, b = torch.empty(N, 324, device=dev), torch.empty(N, 324, device=dev)
        ops.corr_lookup(lb, hs, ws, coords, a, 324, N)
        ops.corr_lookup([t.float() for t in lb], hs, ws, coords, b, 324, N)
        assert torch.equal(a, b)


def test_corr_pyramid_fused_rejects_bad_args(dev):
    from zero_tig_b200 import ops
    t = torch.zeros(64, device=dev)
    with pytest.raises(RuntimeError):
        ops.corr_pyramid(t, 256, 4, 4, [t, t, t, t], t)          # grid below 8x8
    f = torch.zeros(2 * 64, 256, device=dev)
    with pytest.raises(RuntimeError):
        ops.corr_pyramid(f, 256, 8, 8, [t, t, t, t], t)          # scratch too small


def test_corr_lookup_vs_reference_golden(gold, dev):
    from zero_tig_b200 import ops
    pyr = [torch.from_numpy(gold[f"lookup_pyr{i}"]) for i in range(4)]
    hs = [p.shape[2] for p in pyr]
    ws = [p.shape[3] for p in pyr]
    lv = [p.reshape(6, -1).cuda().contiguous() for p in pyr]
    coords = torch.from_numpy(gold["lookup_coords"])  # [1,2,2,3]
    c = coords.permute(0, 2, 3, 1).reshape(6, 2).cuda().contiguous()
    out = torch.empty(6, 324, device=dev)
    ops.corr_lookup(lv, hs, ws, c, out, 324, 6)
    ref = torch.from_numpy(gold["lookup_out"]).permute(0, 2, 3, 1).reshape(6, 324)
    assert (out.cpu() - ref).abs().max().item() < 5e-6


def test_corr_lookup_random_vs_oracle(dev):
    from zero_tig_b200 import ops
    import zt_oracle as O
    g = torch.Generator().manual_seed(9)
    h, w = 17, 21
    N = h * w
    sizes = [(17, 21), (8, 10), (4, 5), (2, 2)]
    pyr = [torch.randn(N, 1, a, b, generator=g) for a, b in sizes]
    coords = O.coords_grid(1, h, w) + torch.randn(1, 2, h, w, generator=g) * 6   # many out-of-range taps
    ref = O.corr_lookup(pyr, coords).permute(0, 2, 3, 1).reshape(N, 324)
    lv = [p.reshape(N, -1).cuda().contiguous() for p in pyr]
    out = torch.empty(N, 324, device=dev)
    ops.corr_lookup(lv, [s[0] for s in sizes], [s[1] for s in sizes],
                    coords.permute(0, 2, 3, 1).reshape(N, 2).cuda().contiguous(), out, 324, N)
    assert (out.cpu() - ref).abs().max().item() < 1e-5


def test_lookup_refuses_one_pixel_level(dev):
    from zero_tig_b200 import ops
    t = torch.zeros(4, 4, device=dev)
    with pytest.raises(RuntimeError, match="NaN"):
        ops.corr_lookup([t, t, t, t], [2, 2, 2, 1], [2, 2, 2, 1], torch.zeros(4, 2, device=dev),
                        torch.zeros(4, 324, device=dev), 324, 4)


def test_convex_upsample_vs_reference_golden(gold, dev):
    from zero_tig_b200 import ops
    flow = nhwc(torch.from_numpy(gold["up_flow"]), 4)
    mask = nhwc(torch.from_numpy(gold["up_mask"]), 576)
    out = torch.empty(1, 2, 24, 32, device=dev)
    ops.convex_upsample(flow, 4, mask, 576, out, 3, 4)
    np.testing.assert_allclose(out.cpu().numpy(), gold["up_out"], rtol=0, atol=2e-5)


def test_warp_pair_vs_reference_golden(gold, dev):
    from zero_tig_b200 import ops
    flow = torch.from_numpy(gold["warp_flow"]).cuda()
    img = torch.from_numpy(gold["warp_img"]).cuda()
    img_b = (1 - img).contiguous()
    oa, ob = torch.empty_like(img), torch.empty_like(img)
    ops.warp_pair(flow, 24, 32, img, img_b, oa, ob, 3, 70, 96)
    ref = torch.from_numpy(gold["warp_out"])
    # coordinates carry ~1e-5 px of fp32 rounding; image gradients are O(1)/px
    assert (oa.cpu() - ref).abs().max().item() < 2e-4
    assert (oa.cpu() - ref).abs().mean().item() < 2e-6


@pytest.mark.parametrize("shape", [(45, 60, 135, 180), (46, 80, 135, 240), (24, 32, 70, 95)])
def test_warp_pair_random_vs_oracle(shape, dev):
    from zero_tig_b200 import ops
    import zt_oracle as O
    hf, wf, H, W = shape
    g = torch.Generator().manual_seed(H)
    flow = torch.randn(1, 2, hf, wf, generator=g) * 4
    a, b = torch.rand(1, 3, H, W, generator=g), torch.rand(1, 3, H, W, generator=g)
    oa, ob = torch.empty(1, 3, H, W, device=dev), torch.empty(1, 3, H, W, device=dev)
    ops.warp_pair(flow.cuda(), hf, wf, a.cuda(), b.cuda(), oa, ob, 3, H, W)
    for got, src in ((oa, a), (ob, b)):
        ref = O.warp_tensor(flow, src)
        assert (got.cpu() - ref).abs().mean().item() < 5e-6
        assert (got.cpu() - ref).abs().max().item() < 5e-4


def test_equalize_bit_exact_vs_reference_golden(gold, dev):
    from zero_tig_b200 import ops
    for k in ("1", "2"):
        q = torch.from_numpy(gold["eq_in" + k]).cuda().contiguous()
        _, h, w = q.shape
        hist = torch.stack([torch.bincount(q[c].reshape(-1).long(), minlength=256) for c in range(3)]).int().contiguous()
        lut = torch.empty(768, dtype=torch.uint8, device=dev)
        ops.equalize_lut(hist, lut)
        eq = torch.empty_like(q)
        packed = torch.empty(h * w, 4, device=dev)
        ops.equalize_apply_pack(q, lut, h, w, h, w, packed, eq)
        assert np.array_equal(eq.cpu().numpy(), gold["eq_out" + k])
        ref = 2 * (torch.from_numpy(gold["eq_out" + k]).float() / 255.0) - 1.0
        assert torch.equal(packed[:, :3].cpu(), ref.permute(1, 2, 0).reshape(-1, 3))


def test_resize_quant_hist_and_image1_vs_oracle(gold, dev):
    """update_cache pre-processing incl. replicate padding and the uint8 truncation cliff."""
    from zero_tig_b200 import ops
    import zt_oracle as O
    g = torch.Generator().manual_seed(21)
    for (H, W, s) in ((390, 400, 3), (128, 160, 1), (100, 131, 2)):
        prev = torch.rand(1, 3, H, W, generator=g)
        L2 = torch.rand(1, 3, H, W, generator=g) * 0.3
        im1, im2 = O.raft_inputs(prev, L2, s)
        ht, wd = H // s, W // s
        hp, wp = (ht + 7) // 8 * 8, (wd + 7) // 8 * 8
        x2 = torch.empty(2, hp, wp, 4, device=dev)
        q = torch.empty(3, ht, wd, dtype=torch.uint8, device=dev)
        hist = torch.empty(768, dtype=torch.int32, device=dev)
        lut = torch.empty(768, dtype=torch.uint8, device=dev)
        eq = torch.empty(3, ht, wd, dtype=torch.uint8, device=dev)
        ops.resize255_image1(prev.cuda(), H, W, ht, wd, x2[0], hp, wp)
        ops.resize255_quant_hist(L2.cuda(), H, W, ht, wd, q, hist)
        ops.equalize_lut(hist, lut)
        ops.equalize_apply_pack(q, lut, ht, wd, hp, wp, x2[1], eq)
        # integer path: quantised image equals the oracle's except where fp32 rounding of the
        # bilinear sum lands within 1 ulp of an integer (allow <= 0.05 % of pixels, by one level)
        ref_q = (O.resize_bilinear(L2, ht, wd) * 255).to(torch.uint8)[0]
        diff = (q.cpu().int() - ref_q.int()).abs()
        assert diff.max().item() <= 1 and (diff > 0).float().mean().item() < 5e-4
        if diff.max().item() == 0:
            assert torch.equal(eq.cpu(), im2[0].to(torch.uint8))
        ref1 = O.raft_pad(im1)
        ref1 = 2 * (ref1 / 255.0) - 1.0
        got1 = x2[0, :, :, :3].permute(2, 0, 1).cpu()
        assert (got1 - ref1[0]).abs().max().item() < 1e-5
        assert x2[0, :, :, 3].abs().max().item() == 0


def test_resize_vs_reference_golden(gold, dev):
    from zero_tig_b200 import ops
    x = torch.from_numpy(gold["rz_in"]).cuda()
    out = torch.empty(1, 16, 24, 4, device=dev)   # wd 21 -> padded 24
    ops.resize255_image1(x, 50, 64, 16, 21, out, 16, 24)
    got = (out[0, :, 1:22, :3].permute(2, 0, 1).cpu() + 1.0) / 2.0   # undo 2*(v/255)-1, pad left = 3//2 = 1
    np.testing.assert_allclose(got.numpy(), gold["rz_out"][0], rtol=0, atol=1e-6)


def test_retinex_elementwise_and_quantize(dev):
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(2)
    HW = 37 * 53
    x = torch.rand(1, 3, 37, 53, generator=g)
    d = torch.randn(HW, 8, generator=g) * 0.2
    xs, ds = x.cuda(), d.cuda()
    L2 = torch.empty_like(xs)
    ops.retinex_l2(xs, ds, 8, L2, HW, 1e-4)
    dn = d[:, :3].t().reshape(1, 3, 37, 53)
    assert torch.equal(L2.cpu(), torch.clamp((x + 1e-4) - dn, 1e-4, 1))
    s = torch.rand(HW, 4, generator=g)
    s2, H2 = torch.empty_like(xs), torch.empty_like(xs)
    ops.retinex_h2(xs, s.cuda(), 4, s2, H2, HW, 1e-4)
    sn = torch.clamp(s[:, :3].t().reshape(1, 3, 37, 53), 1e-4, 1)
    assert torch.equal(s2.cpu(), sn)
    assert (H2.cpu() - torch.clamp((x + 1e-4) / sn, 1e-4, 1)).abs().max().item() < 1e-6
    H3, s3 = torch.empty_like(xs), torch.empty_like(xs)
    ops.retinex_h5(H2, s2, ds, 8, H3, s3, HW, 1e-4)
    assert torch.equal(H3.cpu(), torch.clamp(H2.cpu() - dn, 1e-4, 1))
    assert torch.equal(s3.cpu(), torch.clamp(s2.cpu() - d[:, 3:6].t().reshape(1, 3, 37, 53), 1e-4, 1))
    u8 = torch.empty(37, 53, 3, dtype=torch.uint8, device=dev)
    ops.quantize_u8_hwc(H2, u8, 3, HW)
    ref = np.clip(H2[0].permute(1, 2, 0).cpu().numpy() * 255, 0, 255).astype(np.uint8)   # predict.py:57-61
    assert np.array_equal(u8.cpu().numpy(), ref)


@pytest.mark.parametrize("cin,cout,k,pad,B,H,W", [(4, 64, 7, 3, 2, 40, 56), (64, 96, 3, 1, 1, 24, 40),
                                                   (96, 128, 3, 1, 2, 16, 24)])
def test_stride2_conv_via_space_to_depth(cin, cout, k, pad, B, H, W):
    """A stride-2 convolution (extractor.py:44-48,131) equals the stride-1 tensor-core convolution over the
    space-to-depth input with ops.s2d_weight's rearranged kernel (cropped to the reference output size)."""
    import torch.nn.functional as F
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(cin + cout)
    x = torch.randn(B, cin, H, W, generator=g).cuda()
    w = (torch.randn(cout, cin, k, k, generator=g) * (1.0 / (cin * k * k) ** 0.5)).cuda()
    b = (torch.randn(cout, generator=g) * 0.1).cuda()
    ref = F.conv2d(x, w, b, stride=2, padding=pad)
    ho, wo = ref.shape[-2:]
    assert (ho, wo) == (H // 2, W // 2)
    xn = x.permute(0, 2, 3, 1).reshape(B * H * W, cin).contiguous()
    xs = torch.empty(B * ho * wo, 4 * cin, device="cuda")
    ops.space_to_depth2(xn, cin, cin, B, H, W, xs)
    ws, npad = ops.s2d_weight(w, pad)
    wt = ops.ConvW(ws, b)
    out = torch.empty(B * ho * wo, cout, device="cuda")
    ops.conv2d(wt, xs, 4 * cin, 4 * cin, B, ho, wo, out, cout, pad=(npad, npad), tc=3, bn=64, crop=(ho, wo))
    want = ref.permute(0, 2, 3, 1).reshape(B * ho * wo, cout)
    assert (out - want).abs().max().item() < 2e-4 * max(1.0, want.abs().max().item())
    # the 1x1 stride-2 shortcut = 1x1 convolution on block (0,0) of the space-to-depth tensor
    w1 = (torch.randn(cout, cin, 1, 1, generator=g) * 0.1).cuda()
    ref1 = F.conv2d(x, w1, None, stride=2).permute(0, 2, 3, 1).reshape(B * ho * wo, cout)
    if cin % 64 == 0 or cin == 96:
        ops.conv2d(ops.ConvW(w1), xs, 4 * cin, cin, B, ho, wo, out, cout, tc=3, bn=64)
        assert (out - ref1).abs().max().item() < 2e-4 * max(1.0, ref1.abs().max().item())


def test_raft_small_kernels_vs_torch():
    """convf1 (7x7, 2->128, ReLU) and the flow head's 3x3 256->2 on the CUDA cores against torch conv2d."""
    import torch.nn.functional as F
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(21)
    H, W = 23, 40
    flow = torch.randn(1, 2, H, W, generator=g).cuda() * 3
    w = (torch.randn(128, 2, 7, 7, generator=g) * 0.1).cuda()
    b = (torch.randn(128, generator=g) * 0.1).cuda()
    ref = torch.relu(F.conv2d(flow, w, b, padding=3))[0].permute(1, 2, 0).reshape(H * W, 128)
    f4 = torch.zeros(H * W, 4, device="cuda")
    f4[:, :2] = flow[0].permute(1, 2, 0).reshape(H * W, 2)
    out = torch.empty(H * W, 128, device="cuda")
    ops.raft_convf1(f4, 4, w.permute(2, 3, 1, 0).contiguous(), b, out, 128, H, W)
    assert (out - ref).abs().max().item() < 1e-4
    x = torch.randn(1, 256, H, W, generator=g).cuda()
    w2 = (torch.randn(2, 256, 3, 3, generator=g) * 0.05).cuda()
    b2 = torch.randn(2, generator=g).cuda()
    ref2 = F.conv2d(x, w2, b2, padding=1)[0].permute(1, 2, 0).reshape(H * W, 2)
    xn = x[0].permute(1, 2, 0).reshape(H * W, 256).contiguous()
    d = torch.zeros(H * W, 4, device="cuda")
    ops.conv3x3_n2(xn, 256, 256, w2.permute(2, 3, 1, 0).reshape(9, 256, 2).contiguous(), b2, d, 4, H, W)
    assert (d[:, :2] - ref2).abs().max().item() < 1e-4 and torch.all(d[:, 2:] == 0)


def test_flow_head2_update_matches_separate_kernels(dev):
    """FlowHead.conv2 fused with coords1 += delta and the two flow outputs == conv3x3_n2 followed by flow_update."""
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(12)
    H, W = 13, 19
    N = H * W
    xn = nhwc(torch.randn(1, 256, H, W, generator=g))
    w9 = (torch.randn(9, 256, 2, generator=g) * 0.05).cuda().contiguous()
    b2 = torch.randn(2, generator=g).cuda()
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    c_a = (torch.stack([xs, ys], -1).float().reshape(N, 2) + torch.randn(N, 2, generator=g)).cuda().contiguous()
    c_b = c_a.clone()
    d = torch.zeros(N, 4, device=dev)
    fa, fb = torch.zeros(N, 4, device=dev), torch.zeros(N, 4, device=dev)
    ha, hb = torch.zeros(N, 384, device=dev), torch.zeros(N, 384, device=dev)
    ops.conv3x3_n2(xn, 256, 256, w9, b2, d, 4, H, W)
    ops.flow_update(c_a, d, 4, fa, 4, ha, 384, 382, H, W)
    ops.flow_head2_update(xn, 256, 256, w9, b2, c_b, fb, 4, hb, 384, 382, H, W)
    assert torch.equal(c_a, c_b) and torch.equal(fa, fb) and torch.equal(ha, hb)
    assert ha[:, 382:].abs().sum().item() > 0


"""GPU: the TMA-fed tcgen05 convolution over bf16 NHWC activations (csrc/conv_tma.cu) against torch
conv2d on the same bf16-rounded operands (fp32 accumulation): all layer shapes of the three Retinex
networks, ragged image sizes (partial tiles, zero padding through TMA out-of-bounds fill)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ACTS = {"none": lambda v: v, "relu": torch.relu, "lrelu": lambda v: F.leaky_relu(v, 0.2), "sigmoid": torch.sigmoid}


def _nhwc_bf16(x, cs):
    _, Cn, H, W = x.shape
    t = torch.zeros(H * W, cs, dtype=torch.bfloat16, device="cuda")
    t[:, :Cn] = x[0].permute(1, 2, 0).reshape(H * W, Cn).to(torch.bfloat16)
    return t


@pytest.mark.parametrize("cin,cs_in,cout,k,act,res,f32,H,W", [
    (3, 8, 48, 3, "lrelu", False, False, 37, 45),      # denoise_1.conv1
    (12, 16, 48, 3, "lrelu", False, False, 16, 8),     # denoise_2.conv1, exactly one tile
    (48, 48, 48, 3, "lrelu", False, False, 50, 70),    # conv2
    (48, 48, 3, 1, "none", False, True, 33, 31),       # denoise_1.conv3 (1x1, fp32 head)
    (48, 48, 6, 1, "none", False, True, 21, 64),       # denoise_2.conv3
    (9, 16, 64, 3, "relu", False, False, 19, 23),      # enhance.in_conv
    (64, 64, 64, 3, "relu", True, False, 64, 72),      # enhance block with residual
    (64, 64, 3, 3, "sigmoid", False, True, 41, 29),    # enhance.out_conv
    (64, 64, 64, 3, "relu", True, False, 270, 480),    # many tiles per CTA (ring wrap, TMEM double buffer)
])
def test_conv_bf16_tma_vs_torch(cin, cs_in, cout, k, act, res, f32, H, W):
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(cin * 1000 + cout * 10 + k)
    x = torch.randn(1, cin, H, W, generator=g).cuda()
    w = (torch.randn(cout, cin, k, k, generator=g) * (1.5 / (cin * k * k) ** 0.5)).cuda()
    b = (torch.randn(cout, generator=g) * 0.1).cuda()
    xb, wb = x.to(torch.bfloat16).float(), w.to(torch.bfloat16).float()
    ref = ACTS[act](F.conv2d(xb, wb, b, padding=k // 2))
    xn = _nhwc_bf16(x, cs_in)
    wt = ops.ConvW(w, b, cin_pad=cs_in)
    rn = None
    if res:
        r = torch.randn(1, cout, H, W, generator=g).cuda()
        rn = _nhwc_bf16(r, cout)
        ref = r.to(torch.bfloat16).float() + ref
    if f32:
        ocs = 4 if cout <= 4 else 8
        out = torch.full((H * W, ocs), 7.0, device="cuda")
        ops.conv2d_bf16(wt, xn, cs_in, H, W, out, ocs, act=act, out_f32=True)
        got = out[:, :cout]
        assert torch.all(out[:, cout:] == 0)
        tol = 2e-5
    else:
        out = torch.zeros(H * W, cout, dtype=torch.bfloat16, device="cuda")
        ops.conv2d_bf16(wt, xn, cs_in, H, W, out, cout, act=act, res=rn, res_cs=cout)
        got = out.float()
        tol = 1.0 / 128          # bf16 output rounding (2^-8 relative), values are O(1)
    want = ref[0].permute(1, 2, 0).reshape(H * W, cout)
    err = (got - want).abs().max().item()
    scale = want.abs().max().item()
    assert err <= tol * max(1.0, scale), (err, scale)
    # second call on the same buffers (tensor-map cache hit) gives the same bits
    out2 = torch.zeros_like(out)
    if f32:
        ops.conv2d_bf16(wt, xn, cs_in, H, W, out2, out.shape[1], act=act, out_f32=True)
    else:
        ops.conv2d_bf16(wt, xn, cs_in, H, W, out2, cout, act=act, res=rn, res_cs=cout)
    assert torch.equal(out, out2)


def test_pack_nhwc_bf16():
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(1)
    a, b, c, d = [torch.rand(1, 3, 13, 17, generator=g).cuda() for _ in range(4)]
    out = torch.empty(13 * 17, 16, dtype=torch.bfloat16, device="cuda")
    ops.pack_nhwc_bf16([a, b, c, d], 3, 0.0, out, 16, 13 * 17)
    want = torch.cat([a, b, c, d], 1)[0].permute(1, 2, 0).reshape(-1, 12).to(torch.bfloat16)
    assert torch.equal(out[:, :12], want) and torch.all(out[:, 12:] == 0)
    out8 = torch.empty(13 * 17, 8, dtype=torch.bfloat16, device="cuda")
    ops.pack_nhwc_bf16([a], 3, 1e-4, out8, 8, 13 * 17)
    assert torch.equal(out8[:, :3], (a + 1e-4)[0].permute(1, 2, 0).reshape(-1, 3).to(torch.bfloat16))
    assert torch.all(out8[:, 3:] == 0)


@pytest.mark.parametrize("nsrc,cs", [(1, 8), (2, 8), (3, 16), (4, 16)])
def test_pack_nhwc_bf16_four_pixel_path(nsrc, cs):
    """H*W % 4 == 0 takes the float4 / four-pixels-per-thread kernel; same result as the reference cat + permute."""
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(2 + nsrc)
    H, W = 24, 52
    srcs = [torch.rand(1, 3, H, W, generator=g).cuda() for _ in range(nsrc)]
    out = torch.full((H * W, cs), 7.0, dtype=torch.bfloat16, device="cuda")
    ops.pack_nhwc_bf16(srcs, 3, 1e-4, out, cs, H * W)
    want = (torch.cat(srcs, 1) + 1e-4)[0].permute(1, 2, 0).reshape(-1, 3 * nsrc).to(torch.bfloat16)
    assert torch.equal(out[:, :3 * nsrc], want) and torch.all(out[:, 3 * nsrc:] == 0)



@pytest.mark.parametrize("hn,H,W", [(3, 37, 45), (6, 50, 70), (6, 16, 8), (3, 270, 480), (8, 33, 40), (5, 20, 24), (1, 17, 9)])
def test_conv_bf16_tma_fused_head(hn, H, W):
    """conv2 (3x3, 48 -> 48, LeakyReLU) with conv3 (1x1, 48 -> 3 / 6) applied in its epilogue (model.py:26-27,42-43)
    against torch on the same bf16-rounded conv2 operands.  The 48-channel activation stays in fp32 registers, so the
    result is compared with the fp32 chain (tighter than the two-launch path, which rounds it to bf16 in between)."""
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(100 + hn)
    x = torch.randn(1, 48, H, W, generator=g).cuda()
    w2 = (torch.randn(48, 48, 3, 3, generator=g) * (1.5 / (48 * 9) ** 0.5)).cuda()
    b2 = (torch.randn(48, generator=g) * 0.1).cuda()
    w3 = (torch.randn(hn, 48, 1, 1, generator=g) * (1.5 / 48 ** 0.5)).cuda()
    b3 = (torch.randn(hn, generator=g) * 0.1).cuda()
    xb, w2b = x.to(torch.bfloat16).float(), w2.to(torch.bfloat16).float()
    mid = F.leaky_relu(F.conv2d(xb.double(), w2b.double(), b2.double(), padding=1), 0.2)
    want = F.conv2d(mid, w3.double(), b3.double())[0].permute(1, 2, 0).reshape(H * W, hn).float()
    xn = _nhwc_bf16(x, 48)
    ocs = 4 if hn <= 4 else 8
    out = torch.full((H * W, ocs), 7.0, device="cuda")
    ops.conv2d_bf16(ops.ConvW(w2, b2, cin_pad=48), xn, 48, H, W, out, ocs, act="lrelu", head=ops.ConvW(w3, b3))
    torch.cuda.synchronize()
    assert torch.all(out[:, hn:] == 0)
    err = (out[:, :hn] - want).abs().max().item()
    assert err <= 3e-5 * max(1.0, want.abs().max().item()), err
    # and the two-launch path agrees to its bf16 intermediate rounding
    a2 = torch.zeros(H * W, 48, dtype=torch.bfloat16, device="cuda")
    out2 = torch.zeros(H * W, ocs, device="cuda")
    ops.conv2d_bf16(ops.ConvW(w2, b2, cin_pad=48), xn, 48, H, W, a2, 48, act="lrelu")
    ops.conv2d_bf16(ops.ConvW(w3, b3, cin_pad=48), a2, 48, H, W, out2, ocs, out_f32=True)
    assert (out - out2).abs().max().item() < 0.05


def test_conv_bf16_tma_fused_head_bad_args():
    from zero_tig_b200 import ops
    x = torch.zeros(64, 48, dtype=torch.bfloat16, device="cuda")
    w2 = ops.ConvW(torch.zeros(48, 48, 3, 3, device="cuda"), torch.zeros(48, device="cuda"), cin_pad=48)
    w3 = ops.ConvW(torch.zeros(6, 48, 1, 1, device="cuda"), torch.zeros(6, device="cuda"))
    out = torch.zeros(64, 4, device="cuda")
    with pytest.raises(RuntimeError, match="fused head"):
        ops.conv2d_bf16(w2, x, 48, 8, 8, out, 4, act="lrelu", head=w3)      # 6 head channels into a 4-wide row

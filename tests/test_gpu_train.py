"""GPU: Network.forward (23 tensors), the 17 loss terms and train-mode BatchNorm statistics against the
fixtures of the unmodified reference (tests/golden/train_*.npz) and the CPU oracle.  Gates: outputs
>= 50 dB PSNR (most are ~1e-5 abs), loss total and every term within 1e-3 relative (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, psnr

pytestmark = pytest.mark.gpu

NAMES = ("L_pred1 L_pred2 L2 s2 s21 s22 H2 H11 H12 H13 s13 H14 s14 H3 s3 H3_pred H4_pred "
         "L_pred1_L_pred2_diff H3_denoised1_H3_denoised2_diff H2_blur H3_blur H3_denoised1 H3_denoised2").split()


def _meta(g):
    m = [int(v) for v in g["meta"]]
    return m[:5] + [m[5] if len(m) > 5 else 3]


# "540x960" = BASELINE.json configs[2] at its real size, of_scale 3: RAFT input 180 rows padded to 184 (Q3), swapped
# warp scales 540/184 vs 960/320 (Q2) - fixture from the unmodified reference (oracle/make_golden_train.py)
@pytest.mark.parametrize("tag", ["rlv", "uw", "540x960"])
def test_network_forward_loss_vs_reference_golden(tag, weights):
    import zt_oracle as O
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    g = np.load(os.path.join(GOLDEN, f"train_{tag}.npz"))
    H, W, of_scale, n, stride, seed = _meta(g)
    torch.set_num_threads(os.cpu_count() or 1)
    net = Network(synth.make_args(dataset="underwater" if tag == "uw" else "RLV", of_scale=of_scale)).cuda()
    net.load_state_dict(weights)
    net.train()
    assert net.is_WB == (tag == "uw")
    clip = synth.synthetic_clip(n, H, W, seed=seed, underwater=(tag == "uw"))
    st = O.TrainState()
    for t in range(n):
        x = clip[t:t + 1].cuda()
        net.is_new_seq = (t == 0)
        outs = net(x)
        assert len(outs) == 23
        for name, v in zip(NAMES, outs):
            ref = torch.from_numpy(g[f"{name}_{t}"])
            got = v[0, :, ::stride, ::stride].cpu()
            assert got.shape == ref.shape, name
            if "diff" in name:
                assert (got != ref).float().mean().item() < 5e-3, (name, t)
            else:
                assert psnr(got, ref) >= 50.0, (name, t, psnr(got, ref))
                assert (got - ref).abs().max().item() < 1e-3, (name, t)
        # loss (and BatchNorm running statistics) from a clean copy of the weights, like the fixture
        net.load_state_dict(weights)
        loss = net._loss(x)
        assert loss.dim() == 0 and loss.is_cuda
        ref_loss = float(g[f"loss_{t}"])
        assert abs(loss.item() - ref_loss) <= 1e-3 * abs(ref_loss), (t, loss.item(), ref_loss)
        bn = net.enhance.conv[1]
        assert (bn.running_mean.cpu() - torch.from_numpy(g[f"bn_mean_{t}"])).abs().max().item() < 1e-4
        assert (bn.running_var.cpu() - torch.from_numpy(g[f"bn_var_{t}"])).abs().max().item() < 1e-4
        assert int(bn.num_batches_tracked) == 7 + 3
        # per-term parity against the oracle restatement on the same frame
        with torch.no_grad():
            o_outs = O.network_forward(clip[t:t + 1], weights, st, t == 0, of_scale, training=True)
            o_terms = O.loss_terms(clip[t:t + 1], o_outs, is_WB=(tag == "uw"))
        st.last_H3, st.last_s3 = o_outs[13], o_outs[14]
        got_terms = net.last_loss_terms.cpu()
        for name, a, b in zip(O.LOSS_NAMES, got_terms.tolist(), [float(v) for v in o_terms]):
            assert abs(a - b) <= 1e-3 * abs(b) + 1e-6, (name, t, a, b)
        # ... and against the 17 weighted terms the UNMODIFIED reference produced (loss.py:46-77, call order)
        for name, a, b in zip(O.LOSS_NAMES, got_terms.tolist(), g[f"terms_{t}"].tolist()):
            assert abs(a - b) <= 1e-3 * abs(b) + 1e-6, (name, t, a, b)
        net.load_state_dict(weights)


def test_network_eval_mode_forward(weights):
    """model.eval() (train.py:138): BatchNorm folds, running statistics stay untouched."""
    import zt_oracle as O
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    net = Network(synth.make_args(of_scale=1)).cuda()
    net.load_state_dict(weights)
    net.eval()
    x = synth.synthetic_clip(1, 128, 160, seed=7)
    net.is_new_seq = True
    outs = net(x.cuda())
    with torch.no_grad():
        ref = O.network_forward(x, weights, O.TrainState(), True, 1, training=False)
    for i in (2, 3, 6, 13, 14, 19, 20):
        assert psnr(outs[i].cpu(), ref[i]) >= 50.0, i
    assert int(net.enhance.conv[1].num_batches_tracked) == 7


def test_training_kernels_vs_oracle():
    """pair_downsampler, TextureDifference, blur and the loss reductions on random planes."""
    import zt_oracle as O
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(3)
    x = torch.rand(1, 3, 46, 58, generator=g)
    a, b = ops.pair_downsample(x.cuda(), 1e-4)
    ra, rb = O.pair_downsampler(x + 1e-4)
    assert (a.cpu() - ra).abs().max().item() < 1e-7 and (b.cpu() - rb).abs().max().item() < 1e-7
    y = torch.rand(1, 3, 46, 58, generator=g)
    td = ops.texture_diff(x.cuda(), (x + 0.02 * y).cuda()).cpu()
    assert (td != O.texture_difference(x, x + 0.02 * y)).float().mean().item() < 2e-3
    assert (ops.blur21(x.cuda()).cpu() - O.blur(x)).abs().max().item() < 2e-6
    acc = torch.zeros(8, dtype=torch.float64, device="cuda")
    ops.loss_tv(x.cuda(), acc, 0)
    tv = 2.0 * (acc[0] / (45 * 58) + acc[1] / (46 * 57))
    assert abs(tv.item() - O.tv_loss(x).item()) < 1e-5 * O.tv_loss(x).item()


# ------------------------------------------------------------------------------------------
# backward pass (train.py:126-131)
# ------------------------------------------------------------------------------------------
def _rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


# The enhancer trunk (in_conv / shared block / BatchNorm) sits behind three train-mode BatchNorm
# backward passes whose mean subtractions cancel most of the incoming gradient: the reference's own fp32
# CPU run and the fp32 CPU oracle already differ by 5e-4 there (1e-7 elsewhere).  Those tensors are gated
# at 1e-2, everything else at the 1e-3 of SURVEY 8d.
_TRUNK = ("enhance.in_conv.0.weight", "enhance.in_conv.0.bias", "enhance.conv.0.weight", "enhance.conv.1.weight",
          "enhance.conv.1.bias")


@pytest.mark.parametrize("tag", ["rlv", "uw", "540x960"])
def test_loss_backward_vs_reference_golden(tag, weights):
    """loss.backward() through Network._loss fills .grad of the 20 trainable tensors; compared with the
    gradients of the unmodified reference on frame 1 of the golden clip (RAFT + warp in the loop)."""
    from zero_tig_b200 import synth
    from zero_tig_b200.engine import TRAIN_PARAMS
    from zero_tig_b200.model.model import Network
    g = np.load(os.path.join(GOLDEN, f"train_{tag}.npz"))
    H, W, of_scale, n, stride, seed = _meta(g)
    net = Network(synth.make_args(dataset="underwater" if tag == "uw" else "RLV", of_scale=of_scale)).cuda()
    net.load_state_dict(weights)
    net.train()
    clip = synth.synthetic_clip(n, H, W, seed=seed, underwater=(tag == "uw"))
    for t in range(n):
        x = clip[t:t + 1].cuda()
        net.is_new_seq = (t == 0)
        net(x)
        net.load_state_dict(weights)
        net.zero_grad()
        loss = net._loss(x)
        assert loss.requires_grad and loss.grad_fn is not None
        loss.backward()
        net.load_state_dict(weights)
    named = dict(net.named_parameters())
    assert all(p.grad is None for k, p in named.items() if k.startswith("raft."))
    wscale = torch.from_numpy(g["grad.enhance.conv.0.weight"]).norm().item()
    for k in TRAIN_PARAMS:
        ref = torch.from_numpy(g["grad." + k])
        got = named[k].grad
        assert got is not None and got.shape == ref.shape, k
        if k == "enhance.conv.0.bias":   # a bias in front of BatchNorm: the true gradient is zero (rounding noise)
            assert (got.cpu() - ref).norm().item() <= 1e-4 * wscale, k
            continue
        assert _rel(got, ref) <= (1e-2 if k in _TRUNK else 1e-3), (k, _rel(got, ref))


def test_backward_kernels_vs_autograd():
    """Every adjoint kernel against torch autograd over the oracle restatement (odd sizes, borders)."""
    import torch.nn.functional as F
    import zt_oracle as O
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(11)
    H, W = 38, 50
    x = torch.rand(1, 3, H, W, generator=g)
    y = torch.rand(1, 3, H, W, generator=g)
    go = torch.randn(1, 3, H, W, generator=g)

    def grad_of(fn, *ins):
        ins = [i.clone().requires_grad_(True) for i in ins]
        return torch.autograd.grad(fn(*ins), ins, allow_unused=True)

    # blur^T (reflect folding at the borders)
    ref, = grad_of(lambda a: (O.blur(a) * go).sum(), x)
    out = torch.zeros_like(x).cuda()
    ops.blur21_bwd(go.cuda(), torch.empty_like(out), out)
    assert _rel(out, ref) < 1e-5
    # pair_downsampler^T
    g1, g2 = torch.randn(1, 3, H // 2, W // 2, generator=g), torch.randn(1, 3, H // 2, W // 2, generator=g)
    ref, = grad_of(lambda a: sum((o * gg).sum() for o, gg in zip(O.pair_downsampler(a), (g1, g2))), x)
    out = torch.zeros_like(x).cuda()
    ops.pair_downsample_bwd(g1.cuda(), g2.cuda(), out)
    assert _rel(out, ref) < 1e-6
    # TV
    ref, = grad_of(lambda a: 1600 * O.tv_loss(a), x)
    out = torch.zeros_like(x).cuda()
    ops.loss_tv_bwd(x.cuda(), 1600 * 2.0 / ((H - 1) * W), 1600 * 2.0 / (H * (W - 1)), out)
    assert _rel(out, ref) < 1e-5
    # SmoothLoss w.r.t. the illumination map
    s = 0.2 + 0.6 * torch.rand(1, 3, H, W, generator=g)
    ref = grad_of(lambda a, b: 5 * O.smooth_loss(a, b), x, s)[1]
    from zero_tig_b200.engine import SMOOTH_OFFSETS
    ycc, acc = torch.empty_like(x).cuda(), torch.zeros(12, dtype=torch.float64, device="cuda")
    ops.loss_smooth(x.cuda(), s.cuda(), ycc, acc, 0)
    out = torch.zeros_like(x).cuda()
    ops.loss_smooth_bwd(ycc, s.cuda(), [5 * 2.0 / ((H - dy) * (W - abs(dx))) for dy, dx in SMOOTH_OFFSETS], out)
    assert _rel(out, ref) < 1e-4
    # illumination targets (RLV and white-balance variants)
    L2 = 0.02 + 0.2 * torch.rand(1, 3, H, W, generator=g)
    for is_WB in (False, True):
        n3 = 3.0 * H * W
        eps = 1e-9
        def illum(b):
            L2d = L2
            if is_WB:
                ef = (0.3 / (L2d.mean(dim=(2, 3)) + eps)).unsqueeze(2).unsqueeze(3)
            else:
                Y = L2d[:, 2] * 0.299 + L2d[:, 1] * 0.587 + L2d[:, 0] * 0.144
                ef = (0.5 / (Y.mean(dim=(1, 2)) + eps)).view(-1, 1, 1, 1).repeat(1, 3, 1, 1)
            ef = torch.clamp(ef, 1, 25)
            ratio = torch.pow(torch.tensor(0.7), -ef) / ef
            nl = torch.clamp(L2d / b, eps, 0.8)
            bright = torch.clamp(torch.pow(L2d * ef, ef) * ratio, eps, 1)
            adj = torch.clamp(L2d * ef, eps, 1)
            return 700 * F.mse_loss(b, bright) + 1000 * F.mse_loss(nl, adj)
        ref, = grad_of(illum, s)
        chansum, acc = torch.zeros(4, dtype=torch.float64, device="cuda"), torch.zeros(2, dtype=torch.float64, device="cuda")
        ops.loss_illum(L2.cuda(), s.cuda(), is_WB, chansum, acc, 0)
        out = torch.zeros_like(x).cuda()
        ops.loss_illum_bwd(L2.cuda(), s.cuda(), is_WB, chansum, 700 * 2 / n3, 1000 * 2 / n3, out)
        assert _rel(out, ref) < 1e-4, is_WB
    # Loss_inter (LocalMean^T, quirk Q7)
    D = (torch.rand(1, 1, H, W, generator=g) > 0.6).float()

    def inter(a, b):
        lm1, lm2 = O._box5_reflect(a), O._box5_reflect(b)
        return 10000 * (F.mse_loss(a, (1 - D) * lm1 + a * D) + F.mse_loss(b, (1 - D) * lm2 + a * D))
    r1, r2 = grad_of(inter, x, y)
    o1, o2 = torch.zeros_like(x).cuda(), torch.zeros_like(x).cuda()
    ops.loss_inter_bwd(x.cuda(), y.cuda(), D.cuda(), 10000 * 2 / (3.0 * H * W), torch.empty_like(o1), torch.empty_like(o1),
                       o1, o2)
    assert _rel(o1, r1) < 1e-5 and _rel(o2, r2) < 1e-5
    # Loss_Var through calculate_local_variance
    r1, r2 = grad_of(lambda a, b: 1000 * F.mse_loss(O.local_variance(a), O.local_variance(b - a)), x, y)
    a1, a2 = torch.empty_like(x).cuda(), torch.empty_like(x).cuda()
    acc = torch.zeros(1, dtype=torch.float64, device="cuda")
    ops.loss_var(x.cuda(), y.cuda(), a1, a2, acc, 0)
    o1, o2 = torch.zeros_like(x).cuda(), torch.zeros_like(x).cuda()
    t = [torch.empty_like(o1) for _ in range(3)]
    ops.loss_var_bwd(x.cuda(), y.cuda(), a1, a2, 1000 * 2 / (3.0 * H * W), t[0], t[1], t[2], o1, o2)
    assert _rel(o1, r1) < 1e-4 and _rel(o2, r2) < 1e-4
    # clamp(a / b) w.r.t. b
    ref = grad_of(lambda a, b: (torch.clamp((a + 1e-4) / b, 1e-4, 1) * go).sum(), x, s)[1]
    out = torch.zeros_like(x).cuda()
    ops.div_clamp_bwd(x.cuda(), 1e-4, s.cuda(), go.cuda(), 1e-4, 1.0, out)
    assert _rel(out, ref) < 1e-5


@pytest.mark.parametrize("cin,cout,k,H,W", [(64, 64, 3, 33, 70), (48, 48, 3, 20, 129), (12, 48, 3, 17, 64),
                                             (4, 48, 3, 9, 31), (64, 3, 3, 21, 40), (48, 6, 1, 16, 65),
                                             (48, 3, 1, 8, 8)])
def test_conv_wgrad_dgrad_vs_torch(cin, cout, k, H, W):
    """zt_conv2d_wgrad_nhwc_f32 and the flipped-weight data gradient against torch autograd."""
    import torch.nn.functional as F
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(cin * 100 + cout)
    x = torch.randn(1, cin, H, W, generator=g).requires_grad_(True)
    w = (torch.randn(cout, cin, k, k, generator=g) * 0.1).requires_grad_(True)
    b = torch.randn(cout, generator=g).requires_grad_(True)
    go = torch.randn(1, cout, H, W, generator=g)
    (F.conv2d(x, w, b, padding=k // 2) * go).sum().backward()
    cin_p, co_p = (cin + 3) // 4 * 4, (cout + 3) // 4 * 4
    xn = torch.zeros(H * W, cin_p).cuda()
    xn[:, :cin] = x.detach()[0].permute(1, 2, 0).reshape(H * W, cin)
    gn = torch.zeros(H * W, co_p).cuda()
    gn[:, :cout] = go[0].permute(1, 2, 0).reshape(H * W, cout)
    fw = ops.ConvW(w.detach().cuda(), b.detach().cuda(), cin_pad=cin_p)
    wg = ops.WGrad(fw, w.shape, "cuda")
    wg.accumulate(xn, cin_p, gn, co_p, H, W, (k // 2, k // 2))
    wg.accumulate(xn, cin_p, gn, co_p, H, W, (k // 2, k // 2))     # accumulates
    dw, db = wg.grads()
    assert _rel(dw, 2 * w.grad) < 1e-5 and _rel(db, 2 * b.grad) < 1e-5
    bw = ops.dgrad_weights(w.detach().cuda(), cin_pad=co_p)
    dx = torch.empty(H * W, cin_p).cuda()
    ops.conv2d(bw, gn, co_p, co_p, 1, H, W, dx, cin_p, pad=(k // 2, k // 2), tc=3)
    ref = x.grad[0].permute(1, 2, 0).reshape(H * W, cin)
    assert _rel(dx[:, :cin], ref) < 1e-4


def test_bn_train_backward_vs_torch():
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(5)
    HW, Cn = 37 * 41, 64
    x = torch.randn(1, Cn, 37, 41, generator=g).requires_grad_(True)
    res = torch.randn(1, Cn, 37, 41, generator=g)
    bn = torch.nn.BatchNorm2d(Cn)
    with torch.no_grad():
        bn.weight.normal_(1.0, 0.2, generator=g)
        bn.bias.normal_(0.0, 0.2, generator=g)
    go = torch.randn(1, Cn, 37, 41, generator=g)
    ((res + torch.relu(bn(x))) * go).sum().backward()
    nhwc = lambda t: t.detach()[0].permute(1, 2, 0).reshape(HW, Cn).contiguous().cuda()
    stats = torch.zeros(2 * Cn, dtype=torch.float64, device="cuda")
    out = torch.empty(HW, Cn).cuda()
    gm, bt = bn.weight.detach().cuda(), bn.bias.detach().cuda()
    ops.bn_train(nhwc(x), Cn, out, Cn, nhwc(res), Cn, stats, gm, bt, None, None, HW, Cn, bn.eps, 0.1)
    dx, dgb = torch.empty(HW, Cn).cuda(), torch.zeros(2, Cn).cuda()
    red = torch.zeros(2 * Cn, dtype=torch.float64, device="cuda")
    ops.bn_train_bwd(nhwc(x), Cn, nhwc(go), Cn, stats, red, gm, bt, HW, Cn, bn.eps, dx, Cn, dgb[0], dgb[1])
    assert _rel(dx, nhwc(x.grad)) < 1e-4
    assert _rel(dgb[0], bn.weight.grad) < 1e-5 and _rel(dgb[1], bn.bias.grad) < 1e-5


def test_clip_adam_vs_torch(weights):
    """ClipAdam (flat buffers, fused clip + Adam kernel) against clip_grad_norm_ + torch.optim.Adam."""
    from zero_tig_b200.optim import ClipAdam
    g = torch.Generator().manual_seed(9)
    shapes = [(64, 9, 3, 3), (64,), (48, 48, 3, 3), (3,)]
    ref = [torch.nn.Parameter((torch.randn(s, generator=g) * 0.05).cuda()) for s in shapes]
    mine = [torch.nn.Parameter(p.detach().clone()) for p in ref]
    opt_ref = torch.optim.Adam(ref, lr=1e-4, betas=(0.9, 0.999), weight_decay=3e-4)
    opt = ClipAdam(mine, lr=1e-4, betas=(0.9, 0.999), weight_decay=3e-4, max_norm=5.0)
    for it in range(5):
        grads = [(torch.randn(s, generator=g) * (3.0 if it % 2 else 0.01)).cuda() for s in shapes]
        for p, q, gr in zip(ref, mine, grads):
            p.grad = gr.clone()
            if it < 3:
                q.grad.copy_(gr)        # in place into the flat views
            else:
                q.grad = gr.clone()     # replaced tensors are gathered by step()
        torch.nn.utils.clip_grad_norm_(ref, 5.0)
        opt_ref.step()
        opt.step()
        for p, q in zip(ref, mine):
            assert (p - q).abs().max().item() < 2e-7, it
    assert all(q.data_ptr() >= opt.flat.data_ptr() for q in mine)


def test_training_step_decreases_loss(weights):
    """A few full steps (forward, CUDA backward, fused clip+Adam) on one frame: the loss goes down and the
    torch-side optimiser path of train.py:126-131 gives the same first update."""
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    from zero_tig_b200.optim import ClipAdam
    x = synth.synthetic_clip(1, 96, 128, seed=5).cuda()
    nets = []
    for _ in range(2):
        net = Network(synth.make_args(of_scale=1)).cuda()
        net.load_state_dict(weights)
        net.train()
        nets.append(net)
    a, b = nets
    opt_a = ClipAdam(a.trainable(), lr=1e-4, weight_decay=3e-4, max_norm=5.0)
    opt_b = torch.optim.Adam(b.parameters(), lr=1e-4, betas=(0.9, 0.999), weight_decay=3e-4)
    losses = []
    for it in range(4):
        a.is_new_seq = True
        opt_a.zero_grad()
        la = a._loss(x)
        la.backward()
        opt_a.step()
        losses.append(la.item())
        if it == 0:
            b.is_new_seq = True
            opt_b.zero_grad()
            lb = b._loss(x)
            lb.backward()
            torch.nn.utils.clip_grad_norm_(b.parameters(), 5)
            opt_b.step()
            assert abs(la.item() - lb.item()) < 1e-6 * abs(lb.item())
            from zero_tig_b200.engine import TRAIN_PARAMS
            for k, p, q in zip(TRAIN_PARAMS, a.trainable(), b.trainable()):
                if k == "enhance.conv.0.bias":
                    continue   # zero true gradient (bias before BatchNorm): Adam turns rounding noise into +-lr
                assert (p - q).abs().max().item() < 1e-6, k
    assert losses[-1] < losses[0], losses

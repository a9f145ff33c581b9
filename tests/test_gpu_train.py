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


@pytest.mark.parametrize("tag", ["rlv", "uw"])
def test_network_forward_loss_vs_reference_golden(tag, weights):
    import zt_oracle as O
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    g = np.load(os.path.join(GOLDEN, f"train_{tag}.npz"))
    H, W, of_scale, n, stride = [int(v) for v in g["meta"]]
    net = Network(synth.make_args(dataset="underwater" if tag == "uw" else "RLV", of_scale=of_scale)).cuda()
    net.load_state_dict(weights)
    net.train()
    assert net.is_WB == (tag == "uw")
    clip = synth.synthetic_clip(n, H, W, seed=3, underwater=(tag == "uw"))
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

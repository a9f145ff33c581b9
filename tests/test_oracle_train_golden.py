"""CPU: the oracle's training-path restatement (Network.forward, 17 loss terms) against fixtures
produced by the unmodified reference (oracle/make_golden_train.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

NAMES = ("L_pred1 L_pred2 L2 s2 s21 s22 H2 H11 H12 H13 s13 H14 s14 H3 s3 H3_pred H4_pred "
         "L_pred1_L_pred2_diff H3_denoised1_H3_denoised2_diff H2_blur H3_blur H3_denoised1 H3_denoised2").split()


@pytest.mark.parametrize("tag", ["rlv", "uw", "540x960"])
def test_network_forward_and_loss_vs_reference(tag, weights):
    """"540x960" is BASELINE.json configs[2] at its real size (of_scale 3: pad 180 -> 184 rows, swapped warp scales)."""
    import zt_oracle as O
    from zero_tig_b200 import synth
    g = np.load(os.path.join(GOLDEN, f"train_{tag}.npz"))
    meta = [int(v) for v in g["meta"]]
    H, W, of_scale, n, stride = meta[:5]
    torch.set_num_threads(os.cpu_count() or 1)
    clip = synth.synthetic_clip(n, H, W, seed=meta[5] if len(meta) > 5 else 3, underwater=(tag == "uw"))
    st = O.TrainState()
    for t in range(n):
        x = clip[t:t + 1]
        with torch.no_grad():
            outs = O.network_forward(x, weights, st, t == 0, of_scale, training=True)
            terms = O.loss_terms(x, outs, is_WB=(tag == "uw"))
        for name, v in zip(NAMES, outs):
            ref = torch.from_numpy(g[f"{name}_{t}"])
            got = v[0, :, ::stride, ::stride]
            if "diff" in name:      # binary maps: a threshold flip on a handful of pixels is rounding noise
                assert (got != ref).float().mean().item() < 2e-3, name
            else:
                assert (got - ref).abs().max().item() < 2e-5, (name, t)
        total = float(sum(terms))
        ref_loss = float(g[f"loss_{t}"])
        assert abs(total - ref_loss) <= 1e-4 * abs(ref_loss), (t, total, ref_loss)
        # each of the 17 weighted terms against the reference's own (spied) values, loss.py:46-77
        for name, a, b in zip(O.LOSS_NAMES, [float(v) for v in terms], g[f"terms_{t}"].tolist()):
            assert abs(a - b) <= 1e-4 * abs(b) + 1e-7, (name, t, a, b)
        st.last_H3, st.last_s3 = outs[13], outs[14]


def test_blur_is_separable():
    import zt_oracle as O
    x = torch.rand(1, 3, 40, 50, generator=torch.Generator().manual_seed(0))
    k = O.blur_taps()
    xp = torch.nn.functional.pad(x, (10, 10, 10, 10), mode="reflect")
    h = torch.nn.functional.conv2d(xp, k.view(1, 1, 1, 21).repeat(3, 1, 1, 1), groups=3)
    v = torch.nn.functional.conv2d(h, k.view(1, 1, 21, 1).repeat(3, 1, 1, 1), groups=3)
    assert (v - O.blur(x)).abs().max().item() < 2e-6

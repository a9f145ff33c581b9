"""Training-path fixtures from the UNMODIFIED reference: Network.forward (23 tensors), the
LossFunction total and the parameter gradients, for an RLV and an underwater (is_WB) model.

    python oracle/make_golden_train.py        (build container only; /root/reference needed)

Writes tests/golden/train_{rlv,uw}.npz (128x160, of_scale 1) and tests/golden/train_540x960.npz (BASELINE.json
configs[2] at its real size: 540x960, of_scale 3 - RAFT input padded 180 -> 184 rows (Q3) and unequal swapped warp
scales 540/184 vs 960/320 (Q2)).  Every fixture also holds the 17 weighted terms of LossFunction.forward
(loss.py:46-77), captured by wrapping the criterion's sub-modules.  Weights / clips come from zero_tig_b200.synth.
TEST INFRASTRUCTURE - never imported by the product.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, import_reference  # noqa: E402

NAMES = ("L_pred1 L_pred2 L2 s2 s21 s22 H2 H11 H12 H13 s13 H14 s14 H3 s3 H3_pred H4_pred "
         "L_pred1_L_pred2_diff H3_denoised1_H3_denoised2_diff H2_blur H3_blur H3_denoised1 H3_denoised2").split()


# weights of the 17 terms in call order (loss.py:46-77): two MSE, smooth, TV, thirteen MSE
TERM_WEIGHTS = (700, 1000, 5, 1600, 1000, 1000, 1000, 1000, 1000, 1000, 1000, 1000, 10000, 1000, 10000, 10000, 1000)


def spy_terms(criterion):
    """Record the value every loss sub-module returns, in call order (LossFunction.forward, loss.py:46-77)."""
    calls = []
    for name in ("_l2_loss", "smooth_loss", "L_TV_loss"):
        mod = getattr(criterion, name)
        orig = mod.forward

        def fwd(*a, _o=orig, **k):
            r = _o(*a, **k)
            calls.append(float(r.detach()))
            return r
        mod.forward = fwd
    return calls


def main():
    torch.set_num_threads(8)
    ref_model, ref_loss, ref_utils, _ = import_reference()
    sys.path.insert(0, ROOT)
    from zero_tig_b200 import synth
    for tag, dataset, H, W, of_scale, stride, seed in (("rlv", "RLV", 128, 160, 1, 2, 3), ("uw", "underwater", 128, 160, 1, 2, 3),
                                                       ("540x960", "RLV", 540, 960, 3, 6, 100)):
        torch.manual_seed(0)
        args = synth.make_args(dataset=dataset, of_scale=of_scale)
        net = ref_model.Network(args)
        sd = synth.synthetic_state_dict(net.state_dict(), seed=1234)
        net.load_state_dict(sd)
        net.train()                      # train.py:116 (RAFT is forced to eval inside update_cache)
        calls = spy_terms(net._criterion)
        clip = synth.synthetic_clip(2, H, W, seed=seed, underwater=(dataset == "underwater"))
        out = {"meta": np.array([H, W, of_scale, 2, stride, seed])}
        for t in range(2):
            net.is_new_seq = (t == 0)
            net.zero_grad()
            x = clip[t:t + 1]
            outs = net(x)                # forward only (fresh BN running-stat update happens here too)
            for n, v in zip(NAMES, outs):
                out[f"{n}_{t}"] = v.detach()[0, :, ::stride, ::stride].numpy().copy()
            # loss + gradients on the same frame from a clean copy of the recurrent/BN state
            net.load_state_dict(sd)
            del calls[:]
            loss = net._loss(x)
            assert len(calls) == 17, len(calls)
            out[f"terms_{t}"] = np.array([w * c for w, c in zip(TERM_WEIGHTS, calls)], dtype=np.float64)
            assert abs(out[f"terms_{t}"].sum() - loss.item()) <= 1e-4 * abs(loss.item())
            loss.backward()
            out[f"loss_{t}"] = np.array(loss.item(), dtype=np.float64)
            if t == 1:
                for k, p in net.named_parameters():
                    if p.grad is not None and not k.startswith("raft."):
                        out["grad." + k] = p.grad.numpy().copy()
            out[f"bn_mean_{t}"] = net.enhance.conv[1].running_mean.numpy().copy()
            out[f"bn_var_{t}"] = net.enhance.conv[1].running_var.numpy().copy()
            net.load_state_dict(sd)      # frames are judged with the synthetic weights, not trained ones
        np.savez_compressed(os.path.join(OUT, f"train_{tag}.npz"), **out)
        print(tag, "loss", [float(out[f"loss_{t}"]) for t in range(2)], "keys", len(out))


if __name__ == "__main__":
    main()

"""Evidence for the gradient gate: how far apart are two fp32 executions of the SAME
reference algorithm on the enhancer-trunk gradients?

The oracle (torch functional restatement of the reference, pinned to the unmodified reference's outputs by
tests/test_oracle_*golden.py) is run twice on frame 1 of the golden training clip: on the CPU (torch CPU fp32) and
on the GPU with stock torch/cuDNN/cuBLAS fp32 kernels, TF32 OFF - i.e. what the unmodified reference would compute
on a CUDA device.  Nothing of zero_tig_b200's CUDA path is involved.  Per trainable tensor it prints the relative
L2 difference of the gradients between the two, next to reference(CPU)-vs-oracle(CPU) from the golden file.
    python tools/grad_conditioning.py [rlv|uw|540x960]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import zt_oracle as O                                  # noqa: E402
from zero_tig_b200 import synth                        # noqa: E402
from zero_tig_b200.engine import TRAIN_PARAMS          # noqa: E402
from zero_tig_b200.model.model import Network          # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def grads_on(device, weights, clip, of_scale, is_WB):
    if device != "cpu":
        torch.set_default_device(device)     # the reference CLIs do the same (train.py:49): factories land on the GPU
    try:
        w = {k: v.to(device) for k, v in weights.items()}
        wreq = {k: (v.clone().requires_grad_(True) if k in TRAIN_PARAMS else v) for k, v in w.items()}
        st = O.TrainState()
        out = None
        for t in range(clip.shape[0]):
            x = clip[t:t + 1].to(device)
            outs = O.network_forward(x, wreq, st, t == 0, of_scale, training=True)
            terms = O.loss_terms(x, outs, is_WB=is_WB)
            loss = sum(terms)
            g = torch.autograd.grad(loss, [wreq[k] for k in TRAIN_PARAMS])
            st.last_H3, st.last_s3 = outs[13].detach(), outs[14].detach()
            out = (float(loss), [v.detach().double().cpu() for v in g], outs[2].detach().cpu(), outs[13].detach().cpu())
        return out
    finally:
        torch.set_default_device("cpu")


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "rlv"
    g = np.load(os.path.join(ROOT, "tests", "golden", f"train_{tag}.npz"))
    meta = [int(v) for v in g["meta"]]
    H, W, of_scale, n, stride = meta[:5]
    seed = meta[5] if len(meta) > 5 else 3
    weights = synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)
    clip = synth.synthetic_clip(n, H, W, seed=seed, underwater=(tag == "uw"))
    torch.set_num_threads(os.cpu_count() or 1)
    lc, gc, L2c, H3c = grads_on("cpu", weights, clip, of_scale, tag == "uw")
    lg, gg, L2g, H3g = grads_on("cuda", weights, clip, of_scale, tag == "uw")
    print("# %s %dx%d of_scale %d, frame 1: oracle fp32 on CPU vs the same torch code on CUDA (TF32 off)" % (tag, H, W, of_scale))
    print("# loss cpu %.6f cuda %.6f reference %.6f; max|L2 diff| %.2e, max|H3 diff| %.2e" % (
        lc, lg, float(g["loss_1"]), (L2c - L2g).abs().max().item(), (H3c - H3g).abs().max().item()))
    print("%-28s %10s %16s %18s %18s" % ("tensor", "|g|", "cuda vs cpu", "cuda vs reference", "cpu vs reference"))
    for k, a, b in zip(TRAIN_PARAMS, gc, gg):
        ref = torch.from_numpy(g["grad." + k]).double()
        rel = lambda u, v: ((u - v).norm() / (v.norm() + 1e-30)).item()      # noqa: E731
        print("%-28s %10.3e %16.2e %18.2e %18.2e" % (k, a.norm().item(), rel(b, a), rel(b, ref), rel(a, ref)))


if __name__ == "__main__":
    main()

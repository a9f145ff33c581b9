"""GPU debug tool: gradients of Network._loss (CUDA backward) vs the reference goldens and vs autograd
over the CPU oracle.  Prints the relative L2 error per trainable tensor.
    python tools/grad_check.py [rlv|uw|540x960]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import zt_oracle as O                                  # noqa: E402  (checker)
from zero_tig_b200 import synth                        # noqa: E402
from zero_tig_b200.engine import TRAIN_PARAMS          # noqa: E402
from zero_tig_b200.model.model import Network          # noqa: E402


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "rlv"
    g = np.load(os.path.join(ROOT, "tests", "golden", f"train_{tag}.npz"))
    meta = [int(v) for v in g["meta"]]
    H, W, of_scale, n, stride = meta[:5]
    seed = meta[5] if len(meta) > 5 else 3
    tmpl = Network(synth.make_args()).state_dict()
    weights = synth.synthetic_state_dict(tmpl, seed=1234)
    net = Network(synth.make_args(dataset="underwater" if tag == "uw" else "RLV", of_scale=of_scale)).cuda()
    net.load_state_dict(weights)
    net.train()
    clip = synth.synthetic_clip(n, H, W, seed=seed, underwater=(tag == "uw"))
    wreq = {k: (v.clone().requires_grad_(True) if k in TRAIN_PARAMS else v) for k, v in weights.items()}
    st = O.TrainState()
    for t in range(n):
        x = clip[t:t + 1].cuda()
        net.is_new_seq = (t == 0)
        net(x)
        net.load_state_dict(weights)
        net.zero_grad()
        loss = net._loss(x)
        loss.backward()
        print("frame", t, "loss", loss.item(), "golden", float(g[f"loss_{t}"]))
        # oracle autograd on the same frame
        outs = O.network_forward(clip[t:t + 1], wreq, st, t == 0, of_scale, training=True)
        terms = O.loss_terms(clip[t:t + 1], outs, is_WB=(tag == "uw"))
        ol = sum(terms)
        ograd = torch.autograd.grad(ol, [wreq[k] for k in TRAIN_PARAMS])
        st.last_H3, st.last_s3 = outs[13].detach(), outs[14].detach()
        named = dict(net.named_parameters())
        for k, og in zip(TRAIN_PARAMS, ograd):
            got = named[k].grad.cpu().double()
            e_o = ((got - og.double()).norm() / (og.double().norm() + 1e-30)).item()
            msg = "%-28s |g| %.3e  vs oracle %.2e" % (k, og.norm().item(), e_o)
            if t == 1:
                ref = torch.from_numpy(g["grad." + k]).double()
                msg += "  vs golden %.2e  (oracle vs golden %.2e)" % (
                    ((got - ref).norm() / ref.norm()).item(), ((og.double() - ref).norm() / ref.norm()).item())
            print(msg)
        net.load_state_dict(weights)


if __name__ == "__main__":
    main()

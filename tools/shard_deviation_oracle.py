"""SURVEY 8e report, CPU side: how far does a frame-range shard (halo frames processed from a fresh
``is_new_seq`` start, outputs discarded) deviate from the UNSHARDED sequential run of the same clip?

    python tools/shard_deviation_oracle.py [H W n_frames cut halo...]

Runs the CPU oracle (the reference algorithm) once over the whole clip and once per halo length from
``cut - halo``; prints PSNR(H2), PSNR(H3) of the shard's frames against the unsharded run by distance from the
cut.  The recurrence is model/model.py:324-339; the halo rule is shard.py.  Tool only (uses oracle/)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import zt_oracle as O  # noqa: E402
from zero_tig_b200 import synth  # noqa: E402
from zero_tig_b200.model.model import Network  # noqa: E402


def psnr(a, b):
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return 200.0 if mse == 0 else 10.0 * torch.log10(torch.tensor(1.0 / mse)).item()


def main():
    a = [int(v) for v in sys.argv[1:]]
    H, W, n, cut = (a + [390, 400, 32, 16])[:4] if len(a) < 4 else a[:4]
    halos = a[4:] or [1, 2, 4, 8]
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)
    clip = synth.synthetic_clip(n, H, W, seed=3)
    st, full = O.InferState(), []
    for t in range(n):
        full.append(O.finetune_forward(clip[t:t + 1], sd, st, t == 0, 3))
    print("unsharded run done: %d frames %dx%d; cut at frame %d" % (n, H, W, cut))
    for halo in halos:
        st = O.InferState()
        rows = []
        for t in range(cut - halo, n):
            out = O.finetune_forward(clip[t:t + 1], sd, st, t == cut - halo, 3)
            if t >= cut:
                rows.append((t - cut, psnr(out[0], full[t][0]), psnr(out[1], full[t][1])))
        print("halo %d: " % halo + "  ".join("+%d: H2 %.1f H3 %.1f" % r for r in rows[:6] + rows[-1:]))
        print("   min over shard: H2 %.1f dB, H3 %.1f dB" % (min(r[1] for r in rows), min(r[2] for r in rows)))


if __name__ == "__main__":
    main()

"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference).

Run in the build container only (the reference tree does not travel to the GPU box):
    python oracle/make_golden.py
Three shims, all outside the reference tree: a stub ``matplotlib`` module, a
``torch.load`` wrapper forcing map_location='cpu', and ``Tensor.cuda`` as a no-op.
Weights / clips come from ``zero_tig_b200.synth`` (deterministic from seeds), so fixtures
hold outputs only.  TEST INFRASTRUCTURE - never imported by the product.
"""
import os
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def import_reference():
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    _load = torch.load
    torch.load = lambda f, map_location=None, **kw: _load(f, map_location="cpu", **kw)
    torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    import model.model as ref_model      # noqa: E402
    import loss as ref_loss              # noqa: E402
    import utils.utils as ref_utils      # noqa: E402
    from model.RAFT.corr import CorrBlock  # noqa: E402
    sys.path.remove(REF)
    return ref_model, ref_loss, ref_utils, CorrBlock


def main():
    torch.set_num_threads(8)
    torch.manual_seed(0)
    ref_model, ref_loss, ref_utils, CorrBlock = import_reference()
    sys.path.insert(0, ROOT)
    from zero_tig_b200 import synth
    os.makedirs(OUT, exist_ok=True)

    # ---- key-set pin: our containers must expose the reference's state_dict layout ----
    args = synth.make_args(of_scale=3)
    net = ref_model.Network(args)
    keys = {k: list(v.shape) for k, v in net.state_dict().items()}
    import json
    with open(os.path.join(OUT, "state_dict_keys.json"), "w") as f:
        json.dump(keys, f, indent=0, sort_keys=True)
    sd = synth.synthetic_state_dict(net.state_dict(), seed=1234)

    # ---- Finetunemodel clips ----
    def run_clip(tag, H, W, of_scale, n_frames, stride, keep=None):
        a = synth.make_args(of_scale=of_scale)
        with tempfile.NamedTemporaryFile(suffix=".pt") as tf:
            torch.save({k: v for k, v in sd.items() if not k.startswith("raft.")}, tf.name)
            a.model_pretrain = tf.name
            m = ref_model.Finetunemodel(a)
        m.raft.load_state_dict({k[5:]: v for k, v in sd.items() if k.startswith("raft.")})
        m.eval()
        clip = synth.synthetic_clip(n_frames, H, W, seed=3)
        out = {}
        flows = []
        orig = m.raft.forward

        def spy(i1, i2, **kw):
            r = orig(i1, i2, **kw)
            flows.append((i1.clone(), i2.clone(), r[0].clone(), r[1].clone()))
            return r
        m.raft.forward = spy
        with torch.no_grad():
            for t in range(n_frames):
                m.is_new_seq = (t == 0)
                H2, H3, s3 = m(clip[t:t + 1])
                if keep is not None and t not in keep:
                    continue
                out[f"H2_{t}"] = H2[0, :, ::stride, ::stride].numpy().copy()
                out[f"H3_{t}"] = H3[0, :, ::stride, ::stride].numpy().copy()
                out[f"s3_{t}"] = s3[0, :, ::stride, ::stride].numpy().copy()
        i1, i2, flow_low, flow_up = flows[-1]
        out["raft_im2_last"] = i2[0].numpy().astype(np.uint8)
        out["raft_im1_last"] = i1[0, :, ::stride, ::stride].numpy()
        out["flow_low_last"] = flow_low[0].numpy()
        out["flow_up_last"] = flow_up[0, :, ::stride, ::stride].numpy()
        out["meta"] = np.array([H, W, of_scale, n_frames, stride])
        np.savez_compressed(os.path.join(OUT, f"infer_{tag}.npz"), **out)
        print(tag, {k: v.shape for k, v in out.items() if k != "meta"}, "flow|max|", float(flow_up.abs().max()))

    run_clip("s1_128x160", 128, 160, 1, 3, 1)
    run_clip("s3_390x400", 390, 400, 3, 3, 3)
    # BASELINE.md gate "per frame, through frame 63": the recurrence of model.py:324-339 over a 64-frame clip
    run_clip("s3_390x400_64", 390, 400, 3, 64, 3, keep=(0, 1, 7, 15, 31, 47, 62, 63))

    # ---- per-op vectors ----
    g = torch.Generator().manual_seed(11)
    ops = {}
    # corr lookup (corr.py:29-50): N=6 positions on a 2x3 grid, levels 9x8,4x4,2x2,2x2 (synthetic pyramid)
    cb = CorrBlock.__new__(CorrBlock)
    cb.num_levels, cb.radius = 4, 4
    cb.corr_pyramid = [torch.randn(6, 1, h, w, generator=g) for h, w in ((9, 8), (4, 4), (2, 2), (2, 2))]
    coords = torch.rand(1, 2, 2, 3, generator=g) * 10 - 1.5
    ops["lookup_coords"] = coords.numpy()
    for i, p in enumerate(cb.corr_pyramid):
        ops[f"lookup_pyr{i}"] = p.numpy()
    ops["lookup_out"] = cb(coords).numpy()
    # all-pairs volume + pooling (corr.py:13-27,52-60)
    f1, f2 = torch.randn(1, 256, 9, 10, generator=g), torch.randn(1, 256, 9, 10, generator=g)
    ops["corr_f1"], ops["corr_f2"] = f1.numpy(), f2.numpy()
    full = CorrBlock(f1, f2)
    for i, p in enumerate(full.corr_pyramid):
        ops[f"corr_pyr{i}"] = p.numpy()
    # warp_tensor (utils.py:203-230) with unequal scales (Q2) and large flows (zero padding)
    flow = torch.randn(1, 2, 24, 32, generator=g) * 3
    img = torch.rand(1, 3, 70, 96, generator=g)
    ops["warp_flow"], ops["warp_img"] = flow.numpy(), img.numpy()
    ops["warp_out"] = ref_utils.warp_tensor(flow, img, img)[0].numpy()
    # equalize (torchvision, model.py:359): busy image + a near-constant image (step == 0 identity)
    from torchvision.transforms.functional import equalize
    e1 = (torch.rand(3, 40, 50, generator=g) ** 2 * 200).to(torch.uint8)
    e2 = torch.full((3, 8, 8), 17, dtype=torch.uint8)
    e2[0, 0, 0] = 3
    ops["eq_in1"], ops["eq_out1"] = e1.numpy(), equalize(e1).numpy()
    ops["eq_in2"], ops["eq_out2"] = e2.numpy(), equalize(e2).numpy()
    # convex upsample (raft.py:64-75)
    raft = net.raft
    fl, mk = torch.randn(1, 2, 3, 4, generator=g), torch.randn(1, 576, 3, 4, generator=g)
    ops["up_flow"], ops["up_mask"] = fl.numpy(), mk.numpy()
    ops["up_out"] = raft.upsample_flow(fl, mk).numpy()
    # resize used by update_cache (model.py:351): non-integer scale
    rz = torch.rand(1, 3, 50, 64, generator=g)
    ops["rz_in"] = rz.numpy()
    ops["rz_out"] = torch.nn.functional.interpolate(rz, (16, 21), mode="bilinear").numpy()
    np.savez_compressed(os.path.join(OUT, "ops.npz"), **ops)
    print("ops", {k: v.shape for k, v in ops.items()})


if __name__ == "__main__":
    main()

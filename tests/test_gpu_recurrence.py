"""GPU: parity over the LENGTH of the recurrence and across frame-range shard cuts.

BASELINE.md gates every frame "through frame 63": ``H3, s3`` of frame t feed the enhancer and denoise_2 of frame
t+1 through RAFT + warp (model/model.py:324-339), so rounding differences of the bf16-activation Retinex nets
could accumulate.  These tests run the CUDA path beside the fp32 CPU oracle for a whole 64-frame clip (and 16
frames at 1080p) and check the minimum over frames, and they measure what a frame-range shard (shard.py: one halo
frame from a fresh ``is_new_seq`` start) does to the frames after the cut compared with the UNSHARDED run
(SURVEY 8e).  A report is written to gpurun_out/ when that directory exists."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT, epe, psnr

pytestmark = pytest.mark.gpu

EPE_TOL = 0.01     # px, BASELINE.json north_star
PSNR_TOL = 50.0    # dB


def _report(name, lines):
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, name), "w") as f:
            f.write("\n".join(lines) + "\n")
    print("\n".join(lines))


def _model(weights, tmp_path, of_scale=3):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    ck = tmp_path / "synthetic.pt"
    if not ck.exists():
        torch.save(weights, ck)
    m = Finetunemodel(synth.make_args(of_scale=of_scale, model_pretrain=str(ck))).cuda().eval()
    m.raft.load_state_dict({k[5:]: v for k, v in weights.items() if k.startswith("raft.")})
    return m


def _flow_up(m):
    return [v for k, v in m.raft.engine().ws.bufs.items() if k[0] == "flow_up"][0]


@pytest.mark.parametrize("H,W,n", [(390, 400, 64), (1080, 1920, 16)])
def test_recurrence_vs_oracle_every_frame(H, W, n, weights, tmp_path):
    """min over ALL frames of PSNR(H2), PSNR(H3), PSNR(s3) >= 50 dB against the oracle's own sequential run, flow
    EPE <= 0.01 px on the last pair (and on every 8th pair).  390x400 is the golden clip's size (64 frames, the
    clip length of configs[1]); 1080p x 16 is configs[1]'s frame size."""
    import zt_oracle as O
    from zero_tig_b200 import synth
    torch.set_num_threads(os.cpu_count() or 1)
    m = _model(weights, tmp_path)
    clip = synth.synthetic_clip(n, H, W, seed=3)
    st, tr = O.InferState(), {}
    lines, worst, worst_epe = ["frame  PSNR(H2)  PSNR(H3)  PSNR(s3)  EPE(flow_up)"], 1e9, 0.0
    for t in range(n):
        m.is_new_seq = (t == 0)
        got = m(clip[t:t + 1].cuda())
        ref = O.finetune_forward(clip[t:t + 1], weights, st, t == 0, 3, trace=tr)
        ps = [psnr(v.cpu(), r) for v, r in zip(got, ref)]
        e = epe(_flow_up(m).cpu(), tr["flow_up"]) if t > 0 and (t % 8 == 0 or t == n - 1) else float("nan")
        lines.append("%5d  %8.2f  %8.2f  %8.2f  %s" % (t, ps[0], ps[1], ps[2], "%.5f" % e if e == e else "-"))
        worst = min(worst, *ps)
        if e == e:
            worst_epe = max(worst_epe, e)
    lines.append("min PSNR over %d frames: %.2f dB; max EPE: %.5f px" % (n, worst, worst_epe))
    _report("recurrence_%dx%d_%dframes.txt" % (H, W, n), lines)
    assert worst >= PSNR_TOL, worst
    assert worst_epe <= EPE_TOL, worst_epe


def test_recurrence_64_frames_vs_reference_golden(weights, tmp_path):
    """The same 64-frame clip against frames the UNMODIFIED reference produced (tests/golden/infer_s3_390x400_64.npz,
    frames 0,1,7,15,31,47,62,63 on a stride-3 grid)."""
    from zero_tig_b200 import synth
    g = np.load(os.path.join(GOLDEN, "infer_s3_390x400_64.npz"))
    H, W, of_scale, n, stride = [int(v) for v in g["meta"][:5]]
    m = _model(weights, tmp_path, of_scale)
    clip = synth.synthetic_clip(n, H, W, seed=3).cuda()
    for t in range(n):
        m.is_new_seq = (t == 0)
        H2, H3, s3 = m(clip[t:t + 1])
        if f"H2_{t}" in g:
            for name, v in (("H2", H2), ("H3", H3), ("s3", s3)):
                p = psnr(v[0, :, ::stride, ::stride].cpu(), torch.from_numpy(g[f"{name}_{t}"]))
                assert p >= PSNR_TOL, (name, t, p)
    e = epe(_flow_up(m)[0, :, ::stride, ::stride].cpu(), torch.from_numpy(g["flow_up_last"]))
    assert e <= EPE_TOL, e


def test_shard_cut_deviation_from_unsharded_run(weights, tmp_path):
    """a shard that starts mid-sequence sees a fresh state one frame earlier, so its first frames
    are NOT the unsharded run's.  Measure it: a 48-frame clip as three frame-range shards with ONE halo frame each
    (shard.plan, run interleaved on one GPU by MultiShardPipeline exactly like bench.py) against the oracle's
    unsharded sequential run; PSNR per frame index, gate >= 50 dB on every frame including the first after a cut."""
    import zt_oracle as O
    from zero_tig_b200 import shard, synth
    from zero_tig_b200.pipeline import MultiShardPipeline
    torch.set_num_threads(os.cpu_count() or 1)
    H, W, n, S = 390, 400, 48, 3
    clip = synth.synthetic_clip(n, H, W, seed=3)
    st, full = O.InferState(), []
    for t in range(n):
        full.append(O.finetune_forward(clip[t:t + 1], weights, st, t == 0, 3))
    plans = shard.plan([n], S)
    subs, first = [], []
    for p in plans:
        (w,) = p
        lo = w.halo if w.halo >= 0 else w.start
        subs.append([clip[t:t + 1].pin_memory() for t in range(lo, w.stop)])
        first.append((lo, w.start))
    got = {}

    def emit(s, i, h2, h3):
        lo, start = first[s]
        if lo + i >= start:                      # halo outputs are discarded
            got[lo + i] = (h2.clone(), h3.clone())
    MultiShardPipeline([_model(weights, tmp_path) for _ in range(S)], H, W).run(subs, emit)
    torch.cuda.synchronize()
    assert sorted(got) == list(range(n))
    cuts = [w[0].start for w in plans if w[0].halo >= 0]
    lines = ["3 frame-range shards of a %d-frame %dx%d clip, 1 halo frame each; cuts at frames %s" % (n, H, W, cuts),
             "frame  PSNR(H2)  PSNR(H3)   vs the UNSHARDED oracle run"]
    worst = 1e9
    for t in range(n):
        p2, p3 = psnr(got[t][0], full[t][0]), psnr(got[t][1], full[t][1])
        worst = min(worst, p2, p3)
        mark = "  <- first frame after a cut" if t in cuts else ""
        lines.append("%5d  %8.2f  %8.2f%s" % (t, p2, p3, mark))
    lines.append("min over all frames: %.2f dB" % worst)
    _report("shard_cut_deviation.txt", lines)
    assert worst >= PSNR_TOL, worst

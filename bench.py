#!/usr/bin/env python
"""bench.py - enhanced frames/s of the Zero-TIG hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 1080p|uw4k]

A *step* is one steady-state pass of ``Finetunemodel.forward`` over one frame of a synthetic low-light clip:
denoise_1 -> resize/equalise -> RAFT (12 iterations) -> warp pair -> enhancer -> denoise_2.  ``--config 1080p``
(default) is BASELINE.json configs[1]: a 64-frame 1920x1080 RLV clip per GPU; ``--config uw4k`` is configs[3]: a
3840x2160 underwater clip, frame-sharded across the ranks.  New-sequence / halo frames (no flow) are processed before
the warm-up and are never timed as steps; ``clip_pass`` reports the whole clip including them.

Timing: W warm-up steps, then the timed region of EXACTLY K steps (CUDA events on the launching stream, barrier +
synchronize on both sides, max over ranks) is repeated R times (R chosen so the timed regions add up to >= 2 s);
``value`` is K * world / median region time, min / max are reported beside it.  ``e2e`` runs the same K steps through
the public host-buffer API (``pipeline.MultiShardPipeline``: pinned host frames in, pinned host images out, every
frame's H2D and D2H inside the timed region).

Under torchrun each rank owns its own clip (1080p) or its own frame range of the clip (uw4k): no data-path
collective, weak scaling.  ``--impl reference`` times the CPU oracle port of the reference algorithm
(oracle/zt_oracle.py; the reference itself is pure Python and is absent on the GPU box) on the host cores.
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "enhanced_1080p_frames_per_sec"
UNIT = "frames/s"
OF_SCALE = 3
CONFIGS = {
    # name: (H, W, dataset, frames per rank, BASELINE.json config)
    "1080p": (1080, 1920, "RLV", 64, "configs[1]"),
    "uw4k": (2160, 3840, "underwater", 16, "configs[3]"),
}


_JSON_OUT = None


def claim_stdout():
    """stdout must carry exactly ONE JSON line, but libraries write to file descriptor 1 too (NCCL prints its version
    banner there): keep a private duplicate of fd 1 for the JSON line and point fd 1 at stderr for everything else."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit_line(obj):
    f = _JSON_OUT or sys.stdout
    f.write(json.dumps(obj) + "\n")
    f.flush()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.stop, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    # NVML clocks-event-reason bits (nvml.h): sw_power_cap 0x4, hw_slowdown 0x8, sw_thermal 0x20, hw_thermal 0x40
    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def run_nvml(self):
        """Same four fields through NVML (one library call per sample instead of one nvidia-smi process: eight ranks
        polling nvidia-smi at once got 4 samples in 2.5 s).  Returns False when NVML is unavailable."""
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
            reasons(h)
        except Exception:
            return False
        while not self.stop:
            try:
                r = reasons(h)
                self.rows.append([str(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)), str(mx)] +
                                 ["Active" if r & bit else "Not Active" for _, bit in self.BITS])
            except Exception:
                pass
            time.sleep(0.05)
        return True

    def run(self):
        if os.environ.get("ZT_CLOCKS_NVML", "1") != "0" and self.run_nvml():
            return
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in o.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synthetic_weights():
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    return synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)


def build_model(device, dataset="RLV", sd=None):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    sd = sd if sd is not None else synthetic_weights()
    ck = os.path.join("/tmp", "zt_synthetic_%s_%d.pt" % (dataset, os.getpid()))
    torch.save(sd, ck)
    m = Finetunemodel(synth.make_args(dataset=dataset, of_scale=OF_SCALE, model_pretrain=ck))
    os.remove(ck)
    m.raft.load_state_dict({k[5:]: v for k, v in sd.items() if k.startswith("raft.")})
    return m.to(device).eval(), sd


# ------------------------------------------------------------------------------------------------------------------
# CPU arm
# ------------------------------------------------------------------------------------------------------------------
def workload_string(cfg, extra=""):
    H, W, dataset, n, tag = CONFIGS[cfg]
    return "%s inference %dx%d (BASELINE.json %s), of_scale %d, 12 RAFT iters, steady-state frames%s" % (
        dataset, W, H, tag, OF_SCALE, extra)


def run_reference(args):
    """CPU arm: the oracle port of the reference algorithm on all host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zt_oracle as O
    from zero_tig_b200 import synth
    H, W, dataset, _, _ = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = synthetic_weights()
    budget = float(os.environ.get("ZT_REF_BUDGET_S", "150"))
    clip = synth.synthetic_clip_mt(3, H, W, seed=0, underwater=(dataset == "underwater"))
    st = O.InferState()
    O.finetune_forward(clip[0:1], sd, st, True, OF_SCALE)      # frame 0: new sequence (untimed)
    done, t_total, t0 = 0, 0.0, time.time()
    warm = min(args.warmup, 1)                                  # one warm steady frame, the CPU is slow
    for i in range(warm + args.steps):
        t1 = time.time()
        O.finetune_forward(clip[1 + (i % 2):2 + (i % 2)], sd, st, False, OF_SCALE)
        dt = time.time() - t1
        if i >= warm:
            done += 1
            t_total += dt
        if time.time() - t0 > budget and done >= 1:
            break
    fps = done / t_total
    sample = "%d steady-state frames (of %d requested; %.0f s budget)" % (done, args.steps, budget)
    emit_line({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": warm, "ms_per_step": 1e3 * t_total / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.config), "arm": "CPU oracle port of the reference (oracle/zt_oracle.py)"},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0})


def cpu_baseline(sd, cfg):
    """Oracle port timed on the host cores on a bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zt_oracle as O
    from zero_tig_b200 import synth
    H, W, dataset, _, _ = CONFIGS[cfg]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n = int(os.environ.get("ZT_CPU_BASELINE_FRAMES", "8"))      # ~1.5 s per 1080p frame on 16 cores -> ~12 s of CPU work
    clip = synth.synthetic_clip_mt(n + 1, H, W, seed=0, underwater=(dataset == "underwater"))
    st = O.InferState()
    sdc = {k: v.cpu() for k, v in sd.items()}
    O.finetune_forward(clip[0:1], sdc, st, True, OF_SCALE)
    t0 = time.time()
    done = 0
    for t in range(1, n + 1):
        O.finetune_forward(clip[t:t + 1], sdc, st, False, OF_SCALE)
        done += 1
        if time.time() - t0 > 30.0:
            break
    dt = time.time() - t0
    return {"value": done / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d steady-state frames after 1 new-sequence frame (%.1f s of CPU work)" % (done, dt)}


# ------------------------------------------------------------------------------------------------------------------
# per-kernel profile: every C-ABI call of one steady frame, re-timed by CUDA-graph replay
# ------------------------------------------------------------------------------------------------------------------
class KernelProfiler:
    """Records every ``zero_tig_b200.ops`` call of one eager frame (function + arguments; all buffers are static
    workspaces, so a call can be repeated after the frame), then times each DISTINCT call as REP back-to-back
    launches inside one CUDA graph (CUDA events around the replay on the launching stream): no Python / launch /
    event overhead in the number, programmatic-dependent-launch overlap between the copies as in the real chain.
    HBM-class kernels whose working set would fit the 126 MB L2 are timed with a 256 MB L2 flush between the copies
    (the flush-only graph is subtracted)."""

    SKIP = {"blur_taps", "corr_pyramid_scratch_bytes", "s2d_weight", "dgrad_weights", "check", "lib"}

    def __init__(self, ops_mod, rep=8):
        self.ops, self.rep, self.calls, self.orig = ops_mod, rep, [], {}
        self._flush_buf = None
        self._flush_us = None

    def __enter__(self):
        import types
        for name in dir(self.ops):
            f = getattr(self.ops, name)
            if name.startswith("_") or name in self.SKIP or not isinstance(f, types.FunctionType):
                continue
            if f.__module__ != self.ops.__name__:
                continue
            self.orig[name] = f

            def wrap(*a, _f=f, _n=name, **k):
                r = _f(*a, **k)
                self.calls.append((_n, a, k))
                return r
            setattr(self.ops, name, wrap)
        # the weight-gradient kernel is launched from a method (ops.WGrad.accumulate)
        wg = self.ops.WGrad.accumulate
        self.orig["wgrad"] = wg

        def wrap_wg(obj, *a, **k):
            r = wg(obj, *a, **k)
            self.calls.append(("wgrad", (obj,) + a, k))
            return r
        self.ops.WGrad.accumulate = wrap_wg
        return self

    def __exit__(self, *a):
        for n, f in self.orig.items():
            if n == "wgrad":
                self.ops.WGrad.accumulate = f
            else:
                setattr(self.ops, n, f)

    @staticmethod
    def _key(name, a, k):
        def rep(v):
            if torch.is_tensor(v):
                return ("T", v.data_ptr(), tuple(v.shape), str(v.dtype))
            if isinstance(v, (list, tuple)):
                return tuple(rep(x) for x in v)
            if isinstance(v, (int, float, str, bool, type(None))):
                return v
            return id(v)
        return (name, tuple(rep(v) for v in a), tuple(sorted((kk, rep(v)) for kk, v in k.items())))

    def _time_graph(self, fn, n_inner, flush):
        dev = torch.cuda.current_device()
        side = torch.cuda.Stream(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            fn()
            side.synchronize()
            with torch.cuda.graph(g, stream=side):
                for _ in range(n_inner):
                    if flush:
                        self._flush_buf.zero_()
                    fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (3 * n_inner)       # us per inner iteration

    def time_all(self, flush_pred):
        """Returns a list of dicts {name, args, kwargs, count, us} for the distinct calls, in first-seen order."""
        if self._flush_buf is None:
            self._flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
            self._flush_us = self._time_graph(lambda: None, self.rep, True)
        seen, order = {}, []
        for name, a, k in self.calls:
            key = self._key(name, a, k)
            if key not in seen:
                seen[key] = dict(name=name, args=a, kwargs=k, count=0)
                order.append(key)
            seen[key]["count"] += 1
        out = []
        for key in order:
            c = seen[key]
            f = self.orig[c["name"]]
            fl = bool(flush_pred(c))
            us = self._time_graph(lambda: f(*c["args"], **c["kwargs"]), self.rep, fl)
            c["us"] = max(us - (self._flush_us if fl else 0.0), 0.05)
            c["flushed"] = fl
            out.append(c)
        return out


def classify(c, frame_pixels):
    """family, algorithmic flops, issued tensor flops, algorithmic bytes, label of one recorded ops call."""
    name, a, k = c["name"], c["args"], c["kwargs"]
    if name == "conv2d":
        wt, in0_cs, c0, B, Hn, Wn, out, out_cs = a[0], a[2], a[3], a[4], a[5], a[6], a[7], a[8]
        st = k.get("stride", 1)
        crop = k.get("crop")
        Ho, Wo = crop if crop else (Hn // st, Wn // st)
        M = B * Ho * Wo
        cout = wt.cout
        cin = getattr(wt, "w_raw", None).shape[1] if getattr(wt, "w_raw", None) is not None else wt.cin
        flops = 2.0 * M * cout * cin * wt.kh * wt.kw
        tc = k.get("tc", 0)
        res = 1 if k.get("epi", "plain") in ("add", "add_relu", "gru_q") else 0
        byt = 4.0 * (B * Hn * Wn * cin + M * cout * (1 + res))
        fam = ("conv_tc (full resolution, fp32 activations)" if M >= frame_pixels // 8 else
               "conv_tc RAFT (tcgen05, bf16 hi/lo split, fp32 activations)") if tc else "conv_f32 (FFMA)"
        return fam, flops, flops * (3 if tc == 3 else 1 if tc else 0), byt, "%dx%d %d->%d @%dx%d" % (wt.kh, wt.kw, cin, cout, Ho, Wo)
    if name == "conv2d_bf16":
        wt, cin_s, Hn, Wn, out, ocs = a[0], a[2], a[3], a[4], a[5], a[6]
        cout, cin = wt.w_raw.shape[0], wt.w_raw.shape[1]
        M = Hn * Wn
        head = k.get("head")
        res = k.get("res")
        flops = 2.0 * M * cout * cin * wt.kh * wt.kw + (2.0 * M * cout * head.w_raw.shape[0] if head is not None else 0.0)
        res_b = 2.0 * cout if (res is not None and res.data_ptr() != a[1].data_ptr()) else 0.0
        byt = M * (2.0 * cin_s + (4.0 if (k.get("out_f32") or head is not None) else 2.0) * ocs + res_b)
        lab = "%dx%d %d->%d%s%s" % (wt.kh, wt.kw, cin, cout, " +1x1 head %d" % head.w_raw.shape[0] if head is not None else "",
                                    " +residual" if res is not None else "")
        return "conv_tma (TMA-fed tcgen05, bf16 activations)", flops, flops, byt, lab
    if name == "raft_update":
        h8, w8, iters, weights = a[0], a[1], a[2], a[8]
        N = h8 * w8
        per_iter = sum(2.0 * N * w.w_raw.shape[0] * w.w_raw.shape[1] * w.kh * w.kw for w in weights)   # the 9 tensor-core layers
        cuda_core = 2.0 * N * (128 * 98 + 2 * 2304)                       # convf1 + flow-head conv2 (fp32 FFMA)
        byt = iters * N * (324 * 4 + 4 * 100 * 4 + 8.0)                    # lookup traffic, for reference
        return ("raft_update (persistent TMA-fed tcgen05 update block, all %d iterations)" % iters,
                iters * (per_iter + cuda_core), 3.0 * iters * per_iter, byt, "%d iterations @%dx%d" % (iters, h8, w8))
    if name == "warp_pair":
        flow, hf, wf = a[0], a[1], a[2]
        Cn, Hn, Wn = a[7], a[8], a[9]
        return "warp_pair", 0.0, 0.0, 2.0 * hf * wf * 4 + 4.0 * Cn * Hn * Wn * 4, "flow %dx%d -> %dx%d" % (hf, wf, Hn, Wn)
    if name == "corr_lookup":
        N = a[6]
        return "corr_lookup", 0.0, 0.0, N * (324 * 4 + 4 * 100 * 4 + 8.0), "N=%d" % N
    if name == "corr_pyramid":
        h0, w0, levels = a[2], a[3], a[4]
        N = h0 * w0
        byt = sum(l.numel() * l.element_size() for l in levels) + 2.0 * N * 256 * 4
        return "corr_volume+pyramid", 2.0 * N * N * 256, 0.0, byt, "N=%d" % N
    if name == "wgrad":
        obj, X, x_cs, G, g_cs, Hn, Wn = a[0], a[1], a[2], a[3], a[4], a[5], a[6]
        flops = 2.0 * Hn * Wn * obj.cout * obj.cin * obj.kh * obj.kw
        pad = a[7] if len(a) > 7 else k.get("pad")
        tc = bool(obj.TC and (obj.kh, obj.kw) in ((1, 1), (3, 3)) and tuple(pad) == (obj.kh // 2, obj.kw // 2) and
                  obj.cin <= 64 and obj.cout <= 64)             # the condition of ops.WGrad.accumulate
        fam = ("conv weight gradient (tcgen05, bf16 hi/lo split, K = pixels; incl. the two split + bias-gradient passes)"
               if tc else "conv weight gradient (fp32 FFMA)")
        return (fam, flops, 3.0 * flops if tc else 0.0, 4.0 * Hn * Wn * (x_cs + g_cs),
                "%dx%d %d->%d @%dx%d" % (obj.kh, obj.kw, obj.cin, obj.cout, Hn, Wn))
    if name in ("instnorm",):
        return "instnorm", 0.0, 0.0, 0.0, ""
    if name in ("pack_nhwc_bf16", "pack_nhwc", "retinex_l2", "retinex_h2", "retinex_h5", "quantize_u8_hwc",
                "u8_hwc_to_f32", "resize255_image1", "resize255_quant_hist", "equalize_lut", "equalize_apply_pack"):
        return "frame elementwise (pack / Retinex / resize / equalise)", 0.0, 0.0, 0.0, name
    return "other (" + name + ")", 0.0, 0.0, 0.0, ""


def profile_frame(model, x, pk, frame_pixels):
    """One eager steady-state frame of `model` recorded and re-timed per kernel; returns (families, rows)."""
    from zero_tig_b200 import ops
    eng = model.raft.engine()
    saved = (eng.USE_GRAPH, eng.PARALLEL)
    eng.USE_GRAPH, eng.PARALLEL = 0, 0
    try:
        with KernelProfiler(ops) as kp:
            model.is_new_seq = False
            model(x)
            torch.cuda.synchronize()
        rows = kp.time_all(lambda c: c["name"] in ("warp_pair", "corr_pyramid", "pack_nhwc_bf16", "retinex_l2", "retinex_h2",
                                                    "retinex_h5"))
    finally:
        eng.USE_GRAPH, eng.PARALLEL = saved
    fams = {}
    for c in rows:
        fam, fl, iss, byt, lab = classify(c, frame_pixels)
        c.update(family=fam, flops=fl, issued=iss, bytes=byt, label=lab)
        f = fams.setdefault(fam, dict(us=0.0, launches=0, flops=0.0, issued=0.0, bytes=0.0))
        f["us"] += c["us"] * c["count"]
        f["launches"] += c["count"]
        f["flops"] += fl * c["count"]
        f["issued"] += iss * c["count"]
        f["bytes"] += byt * c["count"]
    return fams, rows


def roofline_entry(name, f, pk, step_us, bound):
    e = {"kernel": name, "bound": bound, "launches_per_frame": f["launches"], "ms_per_frame": f["us"] / 1e3,
         "avg_launch_us": f["us"] / max(f["launches"], 1), "share_of_step": f["us"] / step_us}
    if bound == "tensor":
        tf = f["flops"] / f["us"] / 1e6
        e.update(achieved=tf, peak=pk["tf_sust"], unit="TFLOP/s", frac=tf / pk["tf_sust"],
                 frac_issued=f["issued"] / f["us"] / 1e6 / pk["tf_sust"],
                 note="achieved = algorithmic flops of the reference layers (2*M*cout*cin*taps, no padding); frac_issued "
                      "counts the MMAs actually issued (3 per product for the bf16 hi/lo split); peak = sustained bf16")
        if f["bytes"]:
            e["hbm_gbs"] = f["bytes"] / f["us"] / 1e3
    else:
        gbs = f["bytes"] / f["us"] / 1e3
        e.update(achieved=gbs, peak=pk["hbm"], unit="GB/s", frac=gbs / pk["hbm"])
    return e


# ------------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist
    from zero_tig_b200 import _lib, shard, synth
    from zero_tig_b200.pipeline import MultiShardPipeline
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        # a collective mismatch must cost two minutes, not the default ten
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    _lib.lib()
    H, W, dataset, n_frames, cfg_tag = CONFIGS[args.config]
    # S frame-range shards of this rank's clip (each: own model replica, recurrent state and stream; round-robin
    # enqueue from this thread) - the unit shard.plan deals to GPUs, applied inside one GPU as well (pipeline.py)
    S = max(1, int(os.environ.get("ZT_SHARDS_PER_GPU", "3" if args.config == "1080p" else "1")))
    sd = synthetic_weights()
    models = [build_model(dev, dataset, sd)[0] for _ in range(S)]
    from zero_tig_b200.pipeline import share_sms
    share_sms(models)            # S > 1: each replica's persistent RAFT kernel gets 148 // S CTAs (throughput mode)
    solo = build_model(dev, dataset, sd)[0] if S > 1 else None      # full-grid replica for the single-sequence number
    K, Wm = args.steps, args.warmup

    # ---- the clip: 1080p = one 64-frame clip per rank; uw4k = ONE clip cut into `world` frame ranges (+ halo) ----
    uw = dataset == "underwater"
    if args.config == "uw4k":
        start = rank * n_frames
        lo = max(start - 1, 0)                                   # halo frame of this rank's range
        clip_u8 = synth.synthetic_clip_mt(start + n_frames - lo, H, W, seed=7, underwater=uw, start=lo, as_u8=True)
        rank_new_seq = (rank == 0)
    else:
        clip_u8 = synth.synthetic_clip_mt(n_frames, H, W, seed=rank, underwater=uw, as_u8=True)
        rank_new_seq = True
    clip_u8 = clip_u8.pin_memory()                               # [n][H][W][3] uint8: what PIL hands to ToTensor
    n_local = clip_u8.shape[0]
    clip = clip_u8.to(dev).permute(0, 3, 1, 2).float().div(255).contiguous()     # device-resident fp32 NCHW frames
    # shard s of this rank: frames [first_s, stop_s) where first_s is its seed frame (frame 0 or a halo frame)
    plans = shard.plan([n_local], S)
    ranges = []
    for p in plans:
        (w,) = p
        ranges.append((w.halo if w.halo >= 0 else w.start, w.stop))
    halo_frames = sum(1 for p in plans if p[0].halo >= 0)
    steady = [list(range(a + 1, b)) for a, b in ranges]          # frames that run the full (flow + warp) path
    streams = [torch.cuda.Stream(dev) for _ in range(S)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    pos = [0] * S

    def step_dev(s, m=None):
        f = steady[s][pos[s] % len(steady[s])]
        pos[s] += 1
        m = m or models[s]
        with torch.cuda.stream(streams[s]):
            m.is_new_seq = False
            return m(clip[f:f + 1])

    def region(shards, k):
        """k steady-state frames round-robin over `shards` shards; device time from the fork to the join."""
        cur = torch.cuda.current_stream(dev)
        barrier()
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        for s in range(shards):
            streams[s].wait_event(e0)
        for i in range(k):
            step_dev(i % shards, solo if (shards == 1 and solo is not None) else None)
        for s in range(shards):
            ev = torch.cuda.Event()
            ev.record(streams[s])
            cur.wait_event(ev)
        e1.record(cur)
        barrier()
        return e0.elapsed_time(e1), _lib.launch_count() - l0

    def reps_for(ms_one, lo=5, hi=40):
        return int(min(hi, max(lo, math.ceil(2000.0 / max(ms_one, 1e-3)))))

    # ---- device-resident throughput ----
    for s in range(S):
        with torch.cuda.stream(streams[s]):
            models[s].is_new_seq = True
            models[s](clip[ranges[s][0]:ranges[s][0] + 1])      # seed frame (new sequence / halo), never timed
    for i in range(Wm * S):
        step_dev(i % S)
    if solo is not None:                                        # the lone-sequence replica (full SM grid)
        with torch.cuda.stream(streams[0]):
            solo.is_new_seq = True
            solo(clip[ranges[0][0]:ranges[0][0] + 1])
        for i in range(Wm):
            step_dev(0, solo)
    first_ms, launches = region(S, K)
    R = reps_for(first_ms)
    if world > 1:
        t = torch.tensor([R], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        R = int(t.item())
    with ClockSampler(local) as cs:
        dev_ms = [region(S, K)[0] for _ in range(R)]
        single_ms = [region(1, K)[0] for _ in range(max(3, R // 4))] if S > 1 else None
    clocks = cs.summary()

    # ---- end to end through the public host-buffer API.  Headline: 8-bit HWC frames in (ToTensor on the device) and
    # the two 8-bit HWC images of save_images out; second number: fp32 frames in / fp32 H2, H3 out (r01's e2e) ----
    def e2e_run(input_u8, quantize_u8, reps):
        pipe = MultiShardPipeline(models, H, W, depth=int(os.environ.get("ZT_PIPE_DEPTH", "2")), quantize_u8=quantize_u8,
                                  input_u8=input_u8)
        counts = [K // S + (1 if s < K % S else 0) for s in range(S)]
        cache = {}

        def host_frame(f):
            if input_u8:
                return clip_u8[f]
            if f not in cache:      # fp32 NCHW, as ToTensor leaves it on the host (only the frames this run touches)
                cache[f] = clip_u8[f].permute(2, 0, 1).float().div(255)[None].contiguous().pin_memory()
            return cache[f]
        host = [[host_frame(steady[s][j % len(steady[s])]) for j in range(counts[s])] for s in range(S)]
        sink = []
        pipe.run([h[:min(Wm, 3)] for h in host], lambda s, i, a, b: None, new_sequence=False)     # warm-up (graph capture)
        pipe.run([h[:min(Wm, 3)] for h in host], lambda s, i, a, b: None, new_sequence=False)
        out = []
        for _ in range(reps):
            del sink[:]
            barrier()
            t0 = time.perf_counter()
            pipe.run(host, lambda s, i, a, b: sink.append((s, i)), new_sequence=False)
            torch.cuda.synchronize()
            out.append((time.perf_counter() - t0) * 1e3)
            assert sorted(sink) == [(s, i) for s in range(S) for i in range(counts[s])]
        barrier()
        return out, pipe

    e2e_ms, pipe = e2e_run(True, True, max(5, R // 2))
    e2e32_ms, pipe32 = e2e_run(False, False, 3)

    # ---- the whole clip from scratch, seed / halo frames included (64-frame clip of configs[1]) ----
    pass_pipe = MultiShardPipeline(models, H, W, quantize_u8=True, input_u8=True)
    pass_host = [[clip_u8[f] for f in range(a, b)] for a, b in ranges]
    pass_ms = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        pass_pipe.run(pass_host, lambda s, i, a, b: None, new_sequence=True)
        torch.cuda.synchronize()
        pass_ms.append((time.perf_counter() - t0) * 1e3)
    barrier()

    def reduce_max(vals):
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()
    dev_ms, e2e_ms, e2e32_ms, pass_ms = reduce_max(dev_ms), reduce_max(e2e_ms), reduce_max(e2e32_ms), reduce_max(pass_ms)
    if single_ms is not None:
        single_ms = reduce_max(single_ms)

    train = None
    if not args.no_train and args.config == "1080p":
        # every rank runs the extra profiled step (it holds the gradient all-reduce); only rank 0 records it
        train = train_step_bench(dev, world, sd, profile=not args.no_profile, record=(rank == 0))

    out = None
    if rank == 0:
        pk = peaks()
        med = statistics.median(dev_ms)
        fps = lambda ms: world * K / (ms / 1e3)      # noqa: E731
        out = {
            "metric": METRIC, "value": fps(med), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": med / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 tensor-core products, f32 accumulate (Retinex nets: bf16 activations; RAFT: bf16 hi/lo split = "
                     "fp32-grade products)", "data": "synthetic",
            "config": {"workload": workload_string(args.config),
                       "shards": "%d frame-range shard(s) of the rank's clip interleaved on %d stream(s) of its GPU" % (S, S),
                       "clip": "%d-frame synthetic clip per rank, %d seed/halo frame(s) untimed" % (n_local, S),
                       "l2": "per-frame working set (>2 GB of activations) exceeds the 126 MB L2; no explicit flush",
                       "sharding": "frame ranges + 1 halo frame per shard and rank, no data-path collective"},
            "timing": {"repetitions": R, "region_ms_median": med, "region_ms_min": min(dev_ms), "region_ms_max": max(dev_ms),
                       "ms_per_step_min": min(dev_ms) / K, "ms_per_step_max": max(dev_ms) / K,
                       "timed_seconds_total": sum(dev_ms) / 1e3,
                       "note": "each repetition = exactly K steps, CUDA events, barrier + synchronize on both sides, max over ranks"},
            "clocks": clocks,
            "e2e": {"value": fps(statistics.median(e2e_ms)), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
                    "d2h_bytes_per_step": pipe.d2h_bytes, "repetitions": len(e2e_ms),
                    "min": fps(max(e2e_ms)), "max": fps(min(e2e_ms)),
                    "api": "zero_tig_b200.pipeline.MultiShardPipeline.run(input_u8=True, quantize_u8=True): pinned host "
                           "8-bit HWC frames in (transforms.ToTensor on the device, bit-identical), the two 8-bit HWC "
                           "images of predict.py:57-61 out (quantised on the device, bit-identical); copies on side streams"},
            "e2e_fp32": {"value": fps(statistics.median(e2e32_ms)), "unit": UNIT, "h2d_bytes_per_step": pipe32.h2d_bytes,
                         "d2h_bytes_per_step": pipe32.d2h_bytes,
                         "api": "same pipeline with fp32 NCHW frames in and fp32 H2 / H3 out (what predict.py moves)"},
            "clip_pass": {"frames": n_local - (0 if rank_new_seq else 1), "halo_frames": halo_frames,
                          "seconds": statistics.median(pass_ms) / 1e3,
                          "value": world * (n_local - (0 if rank_new_seq else 1)) / (statistics.median(pass_ms) / 1e3), "unit": UNIT,
                          "note": "one whole clip per rank from scratch through the e2e pipeline: new-sequence frame, halo "
                                  "frames of the shards and all copies included; frames counted once (halo outputs discarded)"},
            "shards_per_gpu": S,
            "single_shard": None if single_ms is None else {
                "value": fps(statistics.median(single_ms)), "unit": UNIT, "ms_per_step": statistics.median(single_ms) / K,
                "note": "one sequence alone on the GPU = the strictly sequential per-clip latency path (the persistent RAFT "
                        "kernel on the full SM grid); same K frames, device-timed"},
            "gpu_launches": int(launches),
        }
        if not args.no_profile:
            step_us = (statistics.median(single_ms) if single_ms else med) / K * 1e3     # one chain alone
            fams, rows = profile_frame(solo or models[0], clip[steady[0][0]:steady[0][0] + 1], pk, H * W)
            order = sorted(fams.items(), key=lambda kv: -kv[1]["us"])
            tensor_f = [(n, f) for n, f in order if f["issued"] > 0]
            hbm_f = [(n, f) for n, f in order if f["issued"] == 0 and f["bytes"] > 0]
            top_name, top = order[0]
            bound = "tensor" if top["issued"] > 0 else "hbm"
            roof = roofline_entry(top_name, top, pk, step_us, bound)
            roof["peak_source"] = pk["src"] + (" sustained bf16 throughput" if bound == "tensor" else " copy bandwidth")
            roof["traffic"] = TRAFFIC.get(top_name) if args.config == "1080p" else None      # captured at 1080p only
            roof["timing"] = ("every distinct launch of one steady frame re-timed as %d back-to-back copies in a CUDA graph "
                              "(CUDA events around the replay); shares are relative to one frame chain alone" % 8)
            out["roofline"] = roof
            out["tensor_kernels"] = [roofline_entry(n, f, pk, step_us, "tensor") for n, f in tensor_f]
            out["hbm_kernels"] = [roofline_entry(n, f, pk, step_us, "hbm") for n, f in hbm_f]
            out["other_kernels"] = [{"kernel": n, "launches_per_frame": f["launches"], "ms_per_frame": f["us"] / 1e3,
                                     "share_of_step": f["us"] / step_us} for n, f in order if f["issued"] == 0 and f["bytes"] == 0]
            out["kernel_time_sum_ms"] = sum(f["us"] for _, f in order) / 1e3
            out["conv_tma_layers"] = [
                {"layer": c["label"], "launches": c["count"], "us": c["us"], "tflops": c["flops"] / c["us"] / 1e6,
                 "frac_tensor": c["flops"] / c["us"] / 1e6 / pk["tf_sust"], "algorithmic_mb": c["bytes"] / 1e6,
                 "gbs": c["bytes"] / c["us"] / 1e3, "frac_hbm": c["bytes"] / c["us"] / 1e3 / pk["hbm"]}
                for c in rows if c["name"] == "conv2d_bf16"]
        if train is not None:
            out["train_step"] = train
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(sd, args.config)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        emit_line(out)


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the family's dominant kernel from the committed
# `ncu --set full` capture (profiles/), keyed by family name; None until captured for the current kernel
TRAFFIC = {
    # profiles/r02_ncu_raft_update.txt: dram__bytes_read.sum + dram__bytes_write.sum of one launch (1080p, 12 iterations)
    "raft_update (persistent TMA-fed tcgen05 update block, all 12 iterations)": 78.09e6 + 35.70e6,
    # profiles/r02_ncu_conv_tma.txt is per layer; the family has no single per-launch figure
}


TRAIN_H, TRAIN_W = 540, 960


def train_step_bench(dev, world, sd, steps=8, warmup=2, profile=False, record=True):
    """configs[2]: zero-shot training steps (Network._loss forward incl. RAFT + warp, CUDA backward, one
    all-reduce of the flat gradient buffer when world > 1, fused clip + Adam) on a synthetic 960x540 clip,
    one clip per rank.  The loss of the first two frames (before any optimiser step) is checked against the value
    the UNMODIFIED reference produced on the same clip and weights (tests/golden/train_540x960.npz)."""
    import numpy as np
    import torch.distributed as dist
    from zero_tig_b200 import _lib, synth
    from zero_tig_b200.model.model import Network
    from zero_tig_b200.optim import ClipAdam
    rank = int(os.environ.get("RANK", "0"))
    gold = np.load(os.path.join(ROOT, "tests", "golden", "train_540x960.npz"))
    gseed = int(gold["meta"][5])
    net = Network(synth.make_args(of_scale=OF_SCALE)).to(dev)
    net.load_state_dict(sd)
    net.train()
    check = []
    gclip = synth.synthetic_clip(2, TRAIN_H, TRAIN_W, seed=gseed).to(dev)
    for t in range(2):                      # parity of the timed workload: same frames / weights as the reference fixture
        net.is_new_seq = (t == 0)
        net(gclip[t:t + 1])
        net.load_state_dict(sd)
        with torch.no_grad():
            lv = float(net._loss(gclip[t:t + 1]))
        net.load_state_dict(sd)
        ref = float(gold["loss_%d" % t])
        check.append({"frame": t, "loss": lv, "reference_loss": ref, "rel_err": abs(lv - ref) / abs(ref)})
        if abs(lv - ref) > 1e-3 * abs(ref):
            raise SystemExit("bench.py: training loss %.6f differs from the reference's %.6f by more than 1e-3 (frame %d)"
                             % (lv, ref, t))
    opt = ClipAdam(net.trainable(), lr=1e-4, weight_decay=3e-4, max_norm=5.0)
    clip = synth.synthetic_clip_mt(warmup + steps + 1, TRAIN_H, TRAIN_W, seed=100 + rank).to(dev)

    def step(i):
        net.is_new_seq = (i == 0)
        opt.zero_grad()
        loss = net._loss(clip[i:i + 1])
        loss.backward()
        opt.step()
        return loss
    for i in range(warmup + 1):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(warmup + 1 + i)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    res = {"workload": "zero-shot training step 960x540 (configs[2]): _loss fwd (RAFT + warp + nets + 17 loss terms), CUDA "
                       "backward, flat-gradient all-reduce (NCCL, world>1), fused clip+Adam; one clip per rank",
           "value": world * 1e3 / ms, "unit": "frames/s", "ms_per_step": ms,
           "gpu_launches_per_step": int((_lib.launch_count() - l0) // steps), "steps": steps, "warmup": warmup + 1,
           "last_loss": float(loss.detach()), "loss_check_vs_reference": check,
           "collective": "1 x all_reduce(sum) of 92620 fp32 per step" if world > 1 else "none (1 GPU)"}
    if profile and not record:
        step(warmup + steps)                 # keeps the collective sequence identical on every rank
        torch.cuda.synchronize()
    if profile and record:
        from zero_tig_b200 import ops
        pk = peaks()
        with KernelProfiler(ops) as kp:
            step(warmup + steps)
            torch.cuda.synchronize()
        rows = kp.time_all(lambda c: False)
        fams = {}
        for c in rows:
            fam, fl, iss, byt, lab = classify(c, TRAIN_H * TRAIN_W)
            if c["name"] == "conv2d":
                fam = "conv_tc training (forward + data gradient, bf16 hi/lo split)" if c["kwargs"].get("tc") else fam
            f = fams.setdefault(fam, dict(us=0.0, launches=0, flops=0.0, issued=0.0, bytes=0.0))
            f["us"] += c["us"] * c["count"]
            f["launches"] += c["count"]
            f["flops"] += fl * c["count"]
            f["issued"] += iss * c["count"]
            f["bytes"] += byt * c["count"]
        order = sorted(fams.items(), key=lambda kv: -kv[1]["us"])
        top_name, top = order[0]
        is_tensor = top["issued"] > 0 or "weight gradient" in top_name      # contractions are judged against the tensor peak
        res["roofline"] = roofline_entry(top_name, top, pk, ms * 1e3, "tensor" if is_tensor else "hbm")
        res["kernel_families"] = [dict(roofline_entry(n, f, pk, ms * 1e3, "tensor") if (f["issued"] > 0 or f["flops"] > 0) else
                                       {"kernel": n, "launches_per_frame": f["launches"], "ms_per_frame": f["us"] / 1e3,
                                        "share_of_step": f["us"] / (ms * 1e3)}) for n, f in order[:10]]
        res["kernel_time_sum_ms"] = sum(f["us"] for _, f in order) / 1e3
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="1080p", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the configs[2] training-step measurement")
    ap.add_argument("--no-profile", action="store_true", help="skip the per-kernel roofline pass")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

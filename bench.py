#!/usr/bin/env python
"""bench.py - enhanced 1080p frames/s of the Zero-TIG hot path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A *step* is one steady-state pass of ``Finetunemodel.forward`` over one 1920x1080 frame of a
synthetic low-light clip (configs[1]): denoise_1 -> resize/equalise -> RAFT (12 iterations)
-> warp pair -> enhancer -> denoise_2.  Frame 0 (new sequence, no flow) is processed before
the warm-up and never timed.  ``value`` is device-timed with the clip resident in HBM;
``e2e`` goes through the public module call with pinned HOST frames (H2D of the frame and
D2H of H2 and H3, the two tensors predict.py:58 copies back) inside the timed region.
Under torchrun each rank owns its own clip shard (frame range + one halo frame): no
data-path collective, weak scaling; timing is barrier-bracketed and max-reduced.

``--impl reference`` times the CPU oracle port of the reference algorithm
(oracle/zt_oracle.py; the reference itself is pure Python and is absent on the GPU box) on
the host cores, one steady-state 1080p frame per step, bounded by a time budget.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "enhanced_1080p_frames_per_sec"
UNIT = "frames/s"
H, W, OF_SCALE = 1080, 1920, 3


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.stop, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in o.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

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


def build_model(device):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel, Network
    tmpl = Network(synth.make_args()).state_dict()
    sd = synth.synthetic_state_dict(tmpl, seed=1234)
    ck = os.path.join("/tmp", "zt_synthetic_RLV_%d.pt" % os.getpid())
    torch.save(sd, ck)
    m = Finetunemodel(synth.make_args(of_scale=OF_SCALE, model_pretrain=ck))
    os.remove(ck)
    m.raft.load_state_dict({k[5:]: v for k, v in sd.items() if k.startswith("raft.")})
    return m.to(device).eval(), sd


# analytic per-frame work at 1080p / of_scale 3 (DESIGN.md "algorithmic work")
CONV_GFLOP_PER_FRAME = 92.0 + 487.3 + 108.7 + 93.9 + 19.3 * 12 + 3.2   # nets + encoders + 12 updates + mask once
WARP_BYTES = 2 * 360 * 640 * 4 + 4 * 3 * H * W * 4                      # flow + 2 planes read + 2 written
LOOKUP_BYTES_PER_ITER = 3600 * (324 * 4 + 4 * 100 * 4 + 8)
# dram__bytes_read.sum + dram__bytes_write.sum of one full-resolution 64->64 conv_tc launch from the
# committed `ncu --set full` capture (profiles/), or None until captured for the current kernel
TRAFFIC_PER_LAUNCH = 483.97e6   # enhancer block launch (64->64 + residual), profiles/r01_ncu_conv_tma_c64.txt


def run_reference(args):
    """CPU arm: oracle port of the reference, all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zt_oracle as O
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Network
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)
    budget = float(os.environ.get("ZT_REF_BUDGET_S", "150"))
    clip = synth.synthetic_clip(3, H, W, seed=0)
    st = O.InferState()
    O.finetune_forward(clip[0:1], sd, st, True, OF_SCALE)      # frame 0: new sequence (untimed)
    done, t_total, t0 = 0, 0.0, time.time()
    warm = min(args.warmup, 1)                                  # one warm steady frame, CPU is slow
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
    sample = "%d steady-state 1080p frames (of %d requested; %.0f s budget)" % (done, args.steps, budget)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": warm, "ms_per_step": 1e3 * t_total / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "RLV inference 1920x1080, of_scale 3, 12 RAFT iters, steady state (CPU oracle port)"},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


class OpTimer:
    """CUDA-event timing of individual C-ABI calls on the launching (current torch) stream;
    used in a separate profiled pass after the headline timing (events perturb the stream)."""

    def __init__(self, ops, names):
        self.ops, self.names, self.rec, self.orig = ops, names, {n: [] for n in names}, {}
        self.bytes, self.flops = {}, {}

    def __enter__(self):
        for n in self.names:
            f = getattr(self.ops, n)
            self.orig[n] = f

            def wrap(*a, _f=f, _n=n, **k):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                r = _f(*a, **k)
                e1.record()
                tag = _n
                if _n == "conv2d":
                    wt, Bn, Hn, Wn = a[0], a[4], a[5], a[6]
                    st = k.get("stride", 1)
                    M = Bn * (Hn // st) * (Wn // st)
                    tag = "corr_volume" if wt.b is None and wt.kh == 1 and wt.cout > 1024 else (
                        "conv_frame" if M >= 500000 else "conv_raft")
                    res = 1 if k.get("epi", "plain") in ("add", "add_relu") else 0
                    self.bytes[tag] = self.bytes.get(tag, 0) + 4.0 * (Bn * Hn * Wn * wt.cin + M * wt.cout * (1 + res))
                    self.flops[tag] = self.flops.get(tag, 0) + 2.0 * M * wt.cout * wt.cin * wt.kh * wt.kw
                if _n == "conv2d_bf16":
                    wt, cin_s, Hn, Wn, ocs = a[0], a[2], a[3], a[4], a[6]
                    tag = "conv_frame"
                    cout, cin_r = wt.w_raw.shape[0], wt.w_raw.shape[1]
                    M = Hn * Wn
                    res = k.get("res")
                    res_b = 2.0 * cout if (res is not None and res.data_ptr() != a[1].data_ptr()) else 0.0   # residual == input: read once
                    byt = M * (2.0 * cin_s + (4.0 if k.get("out_f32") else 2.0) * ocs + res_b)
                    self.bytes[tag] = self.bytes.get(tag, 0) + byt
                    self.flops[tag] = self.flops.get(tag, 0) + 2.0 * M * cout * cin_r * wt.kh * wt.kw
                self.rec.setdefault(tag, []).append((e0, e1))
                return r
            setattr(self.ops, n, wrap)
        return self

    def __exit__(self, *a):
        for n, f in self.orig.items():
            setattr(self.ops, n, f)

    def totals_ms(self):
        torch.cuda.synchronize()
        return {k: (sum(a.elapsed_time(b) for a, b in v), len(v)) for k, v in self.rec.items() if v}


def run_b200(args):
    import torch.distributed as dist
    from zero_tig_b200 import _lib, ops, synth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    # S frame-range shards of the clip per GPU (each: own model replica, recurrent state and stream; round-robin
    # enqueue from this thread) - the unit shard.plan deals to GPUs, applied inside one GPU as well (pipeline.py)
    S = max(1, int(os.environ.get("ZT_SHARDS_PER_GPU", "3")))
    models = []
    for _ in range(S):
        m, sd = build_model(dev)
        models.append(m)
    model = models[0]
    streams = [torch.cuda.Stream(dev) for _ in range(S)]
    K, Wm = args.steps, args.warmup
    n_frames = min(64, (K + S - 1) // S + Wm + 1)
    # each shard of each rank has its own frames (different seed), halo = frame 0
    clips_host = [synth.synthetic_clip(n_frames, H, W, seed=rank * S + s).pin_memory() for s in range(S)]
    clips = [c.to(dev) for c in clips_host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def frame(j):
        return 1 + (j % (n_frames - 1))

    def step_dev(i, shards=S):
        s, j = i % shards, i // shards
        with torch.cuda.stream(streams[s]):
            models[s].is_new_seq = False
            return models[s](clips[s][frame(j):frame(j) + 1])

    def timed_dev(shards):
        """K steady-state frames, round-robin over `shards` shards; device time from the fork to the join."""
        cur = torch.cuda.current_stream(dev)
        barrier()
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(local) as cs:
            e0.record(cur)
            for s in range(shards):
                streams[s].wait_event(e0)
            for i in range(K):
                step_dev(Wm * shards + i, shards)
            for s in range(shards):
                ev = torch.cuda.Event()
                ev.record(streams[s])
                cur.wait_event(ev)
            e1.record(cur)
            barrier()
        return e0.elapsed_time(e1), _lib.launch_count() - l0, cs.summary()

    # ---- device-resident throughput ----
    for s in range(S):
        with torch.cuda.stream(streams[s]):
            models[s].is_new_seq = True
            models[s](clips[s][0:1])      # halo / new-sequence frame, untimed
    for i in range(Wm * S):
        step_dev(i)
    single_ms = None
    if S > 1:
        single_ms, _, _ = timed_dev(1)    # one shard alone (the per-frame latency path), reported beside the headline
    ms, launches, clocks = timed_dev(S)

    # ---- end to end through the public host-buffer API (MultiShardPipeline / FramePipeline: pinned host frames in,
    # pinned host H2/H3 out, every frame's H2D and D2H inside the timed region, overlapped with compute on side
    # streams) ----
    from zero_tig_b200.pipeline import MultiShardPipeline
    pipe = MultiShardPipeline(models, H, W, depth=int(os.environ.get("ZT_PIPE_DEPTH", "2")))
    counts = [K // S + (1 if s < K % S else 0) for s in range(S)]
    host_frames = [[clips_host[s][frame(j):frame(j) + 1] for j in range(counts[s])] for s in range(S)]
    sink = []
    pipe.run([hf[:min(Wm, 3)] for hf in host_frames], lambda s, i, a, b: None, new_sequence=False)
    barrier()
    t0 = time.perf_counter()
    pipe.run(host_frames, lambda s, i, a, b: sink.append((s, i)), new_sequence=False)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert sorted(sink) == [(s, i) for s in range(S) for i in range(counts[s])]
    barrier()

    t = torch.tensor([ms, e2e_s * 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t.tolist()
    train = None
    if not args.no_train:
        tr_ms, tr_launches, tr_loss = train_step_bench(dev, world, sd)
        train = {"workload": "zero-shot training step 960x540 (configs[2]): _loss fwd (RAFT + warp + nets + 17 loss "
                             "terms), CUDA backward, flat-gradient all-reduce (NCCL, world>1), fused clip+Adam; one "
                             "clip per rank", "value": world * 1e3 / tr_ms, "unit": "frames/s", "ms_per_step": tr_ms,
                 "gpu_launches_per_step": int(tr_launches), "steps": 6, "warmup": 3, "last_loss": tr_loss,
                 "collective": "1 x all_reduce(sum) of 92620 fp32 per step" if world > 1 else "none (1 GPU)"}

    out = None
    if rank == 0:
        pk = peaks()
        # ---- per-kernel pass (rank 0): shares + roofline of the dominant kernels ----
        eng = model.raft.engine()
        use_graph, eng.USE_GRAPH = eng.USE_GRAPH, 0     # per-call events need the eager (un-graphed) RAFT pass
        with OpTimer(ops, ["conv2d", "conv2d_bf16", "warp_pair", "corr_lookup", "corr_pyramid", "avgpool2", "instnorm"]) as ot:
            nprof = min(K, 5)
            for i in range(nprof):
                step_dev(i, 1)
            torch.cuda.synchronize()
        eng.USE_GRAPH = use_graph
        tot = ot.totals_ms()
        # shares are relative to ONE frame chain's step time (kernel times of a lone chain / its ms per frame)
        step1_ms = (single_ms if single_ms is not None else ms) / K
        frame_ms = tot["conv_frame"][0] / nprof
        raft_ms = tot["conv_raft"][0] / nprof
        warp_ms = tot["warp_pair"][0] / nprof
        look_ms = tot["corr_lookup"][0] / tot["corr_lookup"][1]
        # fused path: zt_corr_pyramid_f32 (pack + one tcgen05 kernel); ZT_CORR_FUSED=0: 1x1 GEMM + 3 avg-pool kernels
        corr_ms = sum(tot[k][0] for k in ("corr_pyramid", "corr_volume", "avgpool2") if k in tot) / nprof
        # the tcgen05 convolutions of the three full-resolution Retinex nets (conv_tma_kernel: bf16 NHWC
        # activations moved by TMA).  Algorithmic bytes = every input/output/residual activation once;
        # algorithmic flops = 2*M*cout*cin*taps of the reference layers (no padding counted).
        fb = ot.bytes["conv_frame"] / nprof
        gbs = fb / frame_ms / 1e6
        nl = tot["conv_frame"][1] / nprof
        roof = {"bound": "hbm", "kernel": "conv_tma_kernel (TMA-fed tcgen05 implicit-GEMM conv, bf16 activations), the %d "
                                          "full-resolution launches of a frame" % nl,
                "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                "traffic": TRAFFIC_PER_LAUNCH, "algorithmic_bytes_per_launch": fb / nl,
                "peak_source": pk["src"] + " copy bandwidth", "ms_per_frame": frame_ms,
                "avg_launch_ms": frame_ms / nl, "share_of_step": frame_ms / step1_ms,
                "tensor_tflops": ot.flops["conv_frame"] / nprof / frame_ms / 1e9}
        raft_tf = ot.flops["conv_raft"] / nprof / raft_ms / 1e9
        tensor = {"kernel": "conv_tc_kernel + conv_f32_kernel, RAFT launches (M = 3600 positions; bf16 hi/lo split = 3 "
                            "MMAs per useful product)", "bound": "tensor", "achieved": raft_tf, "peak": pk["tf_sust"],
                  "unit": "TFLOP/s (useful fp32-grade)", "frac": raft_tf / pk["tf_sust"], "ms_per_frame": raft_ms,
                  "share_of_step": raft_ms / step1_ms}
        hbm = []
        for name, byt, t_ms in (("warp_pair", WARP_BYTES, warp_ms), ("corr_lookup", LOOKUP_BYTES_PER_ITER, look_ms),
                                ("corr_volume+pyramid", 68.4e6 + 7.4e6, corr_ms)):
            gbs = byt / t_ms / 1e6
            hbm.append({"kernel": name, "bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                        "frac": gbs / pk["hbm"], "ms": t_ms,
                        "timing": "CUDA events around each call in the eager (un-graphed) frame loop, cold caches; "
                                  "includes ~10 us of launch/event overhead per call (ncu kernel times: "
                                  "profiles/r01_launches_200fps.txt, r01_ncu_hbm_kernels.txt)"})
        out = {
            "metric": METRIC, "value": world * K / (ms / 1e3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 tensor-core products, f32 accumulate (Retinex nets: bf16 activations; RAFT: bf16 hi/lo split, f32 storage)", "data": "synthetic",
            "config": {"workload": "RLV inference 1920x1080 (configs[1]), of_scale 3, 12 RAFT iters, steady state, "
                                   "%d frame-range shard(s) per GPU interleaved on %d stream(s) (one halo frame each, "
                                   "untimed), synthetic %d-frame clip per shard" % (S, S, n_frames),
                       "l2": "per-frame working set (>2 GB of activations) exceeds the 126 MB L2; no explicit flush",
                       "sharding": "frame ranges + 1 halo frame per rank, no data-path collective"},
            "clocks": clocks,
            "e2e": {"value": world * K / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
                    "d2h_bytes_per_step": pipe.d2h_bytes,
                    "api": "zero_tig_b200.pipeline.MultiShardPipeline.run (one FramePipeline per shard: pinned host fp32 "
                           "frames in, pinned host fp32 H2/H3 out; copies on side streams)"},
            "shards_per_gpu": S,
            "single_shard": None if single_ms is None else {
                "value": world * K / (single_ms / 1e3), "unit": UNIT, "ms_per_step": single_ms / K,
                "note": "one shard alone on the GPU (per-frame latency path), same K frames, device-timed"},
            "gpu_launches": int(launches), "roofline": roof, "tensor_kernels": [tensor], "hbm_kernels": hbm,
        }
        if train is not None:
            out["train_step"] = train
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(sd)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


TRAIN_H, TRAIN_W = 540, 960


def train_step_bench(dev, world, sd, steps=6, warmup=2):
    """configs[2]: zero-shot training steps (Network._loss forward incl. RAFT + warp, CUDA backward, one
    all-reduce of the flat gradient buffer when world > 1, fused clip + Adam) on a synthetic 960x540 clip,
    one clip per rank.  Returns (ms_per_step max over ranks, launches per step)."""
    import torch.distributed as dist
    from zero_tig_b200 import _lib, synth
    from zero_tig_b200.model.model import Network
    from zero_tig_b200.optim import ClipAdam
    rank = int(os.environ.get("RANK", "0"))
    net = Network(synth.make_args(of_scale=OF_SCALE)).to(dev)
    net.load_state_dict(sd)
    net.train()
    opt = ClipAdam(net.trainable(), lr=1e-4, weight_decay=3e-4, max_norm=5.0)
    clip = synth.synthetic_clip(warmup + steps + 1, TRAIN_H, TRAIN_W, seed=100 + rank).to(dev)

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
    return t.item(), (_lib.launch_count() - l0) // steps, float(loss.detach())


def cpu_baseline(sd):
    """Oracle port timed on the host cores on a bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import zt_oracle as O
    from zero_tig_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n = int(os.environ.get("ZT_CPU_BASELINE_FRAMES", "8"))      # ~1.5 s per 1080p frame on 16 cores -> ~12 s of CPU work
    clip = synth.synthetic_clip(n + 1, H, W, seed=0)
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
            "sample": "%d steady-state 1080p frames after 1 new-sequence frame (%.1f s of CPU work)" % (done, dt)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the configs[2] training-step measurement")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()

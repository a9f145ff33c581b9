"""Host-side orchestration of the sm_100a kernels: RAFT, update_cache, Retinex nets.

The engines read parameters from the (reference-compatible) nn.Module containers,
re-layout them once per device / parameter version, own the NHWC workspaces and issue
the kernel sequence on torch's current stream.  Nothing here computes on the CPU.
"""
import os

import torch

from . import ops
from .ops import ConvW


def _dev_check(t):
    if not t.is_cuda:
        raise RuntimeError("zero_tig_b200 runs on CUDA only (no CPU fallback); got a %s tensor" % t.device)


class _Workspace:
    """Named scratch tensors, allocated on first use per shape and kept (180 GB of HBM)."""

    def __init__(self, device):
        self.device = device
        self.bufs = {}

    def get(self, name, shape, dtype=torch.float32, zero=False):
        key = (name, tuple(shape), dtype)
        t = self.bufs.get(key)
        if t is None:
            t = (torch.zeros if zero else torch.empty)(shape, device=self.device, dtype=dtype)
            self.bufs[key] = t
        return t


# Parameters can be updated through raw pointers (optim.ClipAdam's fused kernel, bn_train's running statistics):
# torch's per-tensor ``_version`` does not see those writes, so every such writer bumps a generation counter of the
# tensors it wrote (keyed by device address) and the packed-weight caches key on it as well.
_PARAM_GENERATION = {}


def bump_param_generation(tensors):
    for t in tensors:
        k = t.data_ptr()
        _PARAM_GENERATION[k] = _PARAM_GENERATION.get(k, 0) + 1


def _param_version(module):
    return tuple((p.data_ptr(), p._version, _PARAM_GENERATION.get(p.data_ptr(), 0))
                 for p in list(module.parameters()) + list(module.buffers()))


# ----------------------------------------------------------------------------------------
# RAFT
# ----------------------------------------------------------------------------------------
def _tc_mode(name, default):
    """Tensor-core mode of a sub-network: 4 = scaled fp16 hi/lo split (three products, ~2^-22 per operand: fp32-grade),
    3 = bf16 hi/lo split (three products, ~2^-17), 1 = plain bf16, 0 = fp32 CUDA-core kernel.  Overridable for
    experiments with ZT_TC_<NAME>."""
    import os
    return int(os.environ.get("ZT_TC_" + name, default))


class RaftEngine:
    """RAFT.forward (model/RAFT/raft.py:77-130) on the kernels in csrc/."""

    TC = _tc_mode("RAFT", 3)
    TC_ENC = _tc_mode("RAFT_ENC", TC)      # fnet / cnet
    TC_UPD = _tc_mode("RAFT_UPD", TC)      # motion encoder + GRU
    TC_HEAD = _tc_mode("RAFT_HEAD", TC)    # flow head + mask head
    TC_CORR = _tc_mode("RAFT_CORR", 3 if TC else 0)   # all-pairs correlation volume
    CORR_FUSED = int(os.environ.get("ZT_CORR_FUSED", "1"))   # volume + pyramid in one kernel (corr_gemm.cu)
    # bf16 STORAGE of the correlation pyramid (configs[4] "fp32 vs bf16 corr"): halves the bytes of the output-bound
    # pyramid kernel; off by default - its EPE against the reference is reported beside the fp32 arm
    CORR_BF16 = int(os.environ.get("ZT_CORR_BF16", "0"))
    BN96 = int(os.environ.get("ZT_ENC_BN96", "1"))
    # N tiles of the two K = 2304 layers of the motion encoder (measured: 48 / 32 instead of 64 put 120 instead of
    # 90 / 60 CTAs on the K loop and change the iteration time by < 1 %: these launches are not K-loop bound)
    BN_C2 = int(os.environ.get("ZT_BN_C2", "96"))
    BN_ENC = int(os.environ.get("ZT_BN_ENC", "64"))
    BN_ZR = int(os.environ.get("ZT_BN_ZR", "128"))      # GRU z|r conv (N = 256)
    BN_Q = int(os.environ.get("ZT_BN_Q", "64"))        # GRU q conv (N = 128)
    BN_C1 = int(os.environ.get("ZT_BN_C1", "128"))      # motion encoder convc1 (N = 256)
    BN_FH = int(os.environ.get("ZT_BN_FH", "128"))      # flow head conv1 (N = 256)
    # all update iterations in one persistent cooperative kernel (csrc/raft_update.cu); 0 = the round-1 launch chain
    UPD_FUSED = int(os.environ.get("ZT_RAFT_FUSED", "1"))
    # N tiles of the nine tensor-core layers inside that kernel (convc1, convf2, convc2, conv, zr1, q1, zr2, q2, fh1)
    FUSED_BN = tuple(int(v) for v in os.environ.get("ZT_RAFT_FUSED_BN", "64,64,64,32,64,32,64,32,64").split(","))
    FUSED_MAX_CTAS = int(os.environ.get("ZT_RAFT_FUSED_CTAS", "0"))

    def __init__(self, raft):
        self.raft = raft
        self._wkey = None
        self.w = None
        self.ws = None

    # ---- weights ----
    def _prepare(self, device):
        key = (str(device), _param_version(self.raft))
        if key == self._wkey:
            return
        sd = {k: v.detach().to(device) for k, v in self.raft.state_dict().items()}
        w = {}

        def bn(prefix):
            return (sd[prefix + ".weight"], sd[prefix + ".bias"], sd[prefix + ".running_mean"],
                    sd[prefix + ".running_var"])

        import torch.nn.functional as F
        for enc, is_bn in (("fnet", False), ("cnet", True)):
            w[f"{enc}.conv1"] = ConvW(sd[f"{enc}.conv1.weight"], sd[f"{enc}.conv1.bias"], cin_pad=4,
                                      bn=bn(f"{enc}.norm1") if is_bn else None)
            # stride-2 convolutions as stride-1 convolutions over the space-to-depth input (tensor-core path)
            w4 = F.pad(sd[f"{enc}.conv1.weight"], (0, 0, 0, 0, 0, 1))
            w[f"{enc}.conv1.s2d"] = ConvW(ops.s2d_weight(w4, 3)[0], sd[f"{enc}.conv1.bias"],
                                          bn=bn(f"{enc}.norm1") if is_bn else None)
            for li in (1, 2, 3):
                for bi in (0, 1):
                    p = f"{enc}.layer{li}.{bi}"
                    w[p + ".conv1"] = ConvW(sd[p + ".conv1.weight"], sd[p + ".conv1.bias"],
                                            bn=bn(p + ".norm1") if is_bn else None)
                    w[p + ".conv2"] = ConvW(sd[p + ".conv2.weight"], sd[p + ".conv2.bias"],
                                            bn=bn(p + ".norm2") if is_bn else None)
                    if p + ".downsample.0.weight" in sd:
                        w[p + ".conv1.s2d"] = ConvW(ops.s2d_weight(sd[p + ".conv1.weight"], 1)[0], sd[p + ".conv1.bias"],
                                                    bn=bn(p + ".norm1") if is_bn else None)
                        w[p + ".down"] = ConvW(sd[p + ".downsample.0.weight"], sd[p + ".downsample.0.bias"],
                                               bn=bn(p + ".norm3") if is_bn else None)
        w["fnet.conv2"] = ConvW(sd["fnet.conv2.weight"], sd["fnet.conv2.bias"])
        w["cnet.conv2.net"] = ConvW(sd["cnet.conv2.weight"], sd["cnet.conv2.bias"], out_slice=slice(0, 128))
        w["cnet.conv2.inp"] = ConvW(sd["cnet.conv2.weight"], sd["cnet.conv2.bias"], out_slice=slice(128, 256))
        u = "update_block."
        for name in ("convc1", "convc2", "convf2", "conv"):
            w["enc." + name] = ConvW(sd[u + f"encoder.{name}.weight"], sd[u + f"encoder.{name}.bias"])
        w["enc.convf1"] = ConvW(sd[u + "encoder.convf1.weight"], sd[u + "encoder.convf1.bias"], cin_pad=4)
        for s in ("1", "2"):
            wz, wr = sd[u + f"gru.convz{s}.weight"], sd[u + f"gru.convr{s}.weight"]
            bz, br = sd[u + f"gru.convz{s}.bias"], sd[u + f"gru.convr{s}.bias"]
            w["gru.zr" + s] = ConvW(torch.cat([wz, wr], 0), torch.cat([bz, br], 0))
            w["gru.q" + s] = ConvW(sd[u + f"gru.convq{s}.weight"], sd[u + f"gru.convq{s}.bias"])
            # the context features `inp` (input channels 128..255 of cat[h, inp, motion], update.py:130) do not
            # change over the 12 iterations: their share of every GRU convolution (+ bias) is computed once per
            # frame and enters the loop convolutions as a per-pixel bias plane; the loop runs on K = 256 channels
            wzr, wq = torch.cat([wz, wr], 0), sd[u + f"gru.convq{s}.weight"]
            w[f"gru.zr{s}.inp"] = ConvW(wzr[:, 128:256].contiguous(), torch.cat([bz, br], 0))
            w[f"gru.q{s}.inp"] = ConvW(wq[:, 128:256].contiguous(), sd[u + f"gru.convq{s}.bias"])
            w[f"gru.zr{s}.hm"] = ConvW(torch.cat([wzr[:, :128], wzr[:, 256:]], 1).contiguous(), None)
            w[f"gru.q{s}.hm"] = ConvW(torch.cat([wq[:, :128], wq[:, 256:]], 1).contiguous(), None)
        # CUDA-core fp32 kernels for the two degenerate GEMMs (K = 98 and N = 2)
        w["enc.convf1.raw"] = (sd[u + "encoder.convf1.weight"].float().permute(2, 3, 1, 0).contiguous(),
                               sd[u + "encoder.convf1.bias"].float().contiguous())
        w["fh2.raw"] = (sd[u + "flow_head.conv2.weight"].float().permute(2, 3, 1, 0).reshape(9, 256, 2).contiguous(),
                        sd[u + "flow_head.conv2.bias"].float().contiguous())
        w["fh1"] = ConvW(sd[u + "flow_head.conv1.weight"], sd[u + "flow_head.conv1.bias"])
        w["fh2"] = ConvW(sd[u + "flow_head.conv2.weight"], sd[u + "flow_head.conv2.bias"])
        w["mask0"] = ConvW(sd[u + "mask.0.weight"], sd[u + "mask.0.bias"])
        w["mask2"] = ConvW(sd[u + "mask.2.weight"], sd[u + "mask.2.bias"])
        self.w, self._wkey = w, key
        if self.ws is None or self.ws.device != device:
            self.ws = _Workspace(device)

    # ---- encoders (extractor.py:167-191) ----
    def _encoder(self, enc, x, B, H, W, is_bn):
        w, ws, T = self.w, self.ws, self.TC_ENC
        stats = ws.get("in_stats", (B * 128 * 2,), torch.float64)
        h, wd = H // 2, W // 2

        def buf(tag, hh, ww, c):
            return ws.get(f"{enc}.{tag}", (B * hh * ww, c))

        y = buf("stem", h, wd, 64)
        s2d = bool(T) and H % 8 == 0 and W % 8 == 0
        if s2d:
            xs = buf("stem.s2d", h, wd, 16)
            ops.space_to_depth2(x, 4, 4, B, H, W, xs)
            ops.conv2d(w[enc + ".conv1.s2d"], xs, 16, 16, B, h, wd, y, 64, pad=(2, 2), act="relu" if is_bn else "none",
                       tc=T, bn=64, crop=(h, wd))
        else:
            ops.conv2d(w[enc + ".conv1"], x, 4, 4, B, H, W, y, 64, stride=2, pad=(3, 3),
                       act="relu" if is_bn else "none")
        if not is_bn:
            ops.instnorm(y, 64, y, 64, stats, B, h * wd, 64, relu_inner=True)
        cin = 64
        for li, (cout, stride) in enumerate(((64, 1), (96, 2), (128, 2)), start=1):
            # one N tile for the 96-channel layers (two tiles of 64 + 32 fetch the A operand twice: §3.7 of DESIGN.md)
            bn_l = 96 if cout == 96 and self.BN96 else 64
            for bi in (0, 1):
                s = stride if bi == 0 else 1
                p = f"{enc}.layer{li}.{bi}"
                ho, wo = h // s if s == 1 else (h + 2 - 3) // s + 1, wd // s if s == 1 else (wd + 2 - 3) // s + 1
                t1 = buf(f"l{li}b{bi}.t1", ho, wo, cout)
                ys = None
                if s != 1 and s2d:
                    ys = buf(f"l{li}.s2d", ho, wo, 4 * cin)
                    ops.space_to_depth2(y, cin, cin, B, h, wd, ys)
                    ops.conv2d(w[p + ".conv1.s2d"], ys, 4 * cin, 4 * cin, B, ho, wo, t1, cout, pad=(1, 1),
                               act="relu" if is_bn else "none", tc=T, bn=bn_l, crop=(ho, wo))
                else:
                    ops.conv2d(w[p + ".conv1"], y, cin, cin, B, h, wd, t1, cout, stride=s, pad=(1, 1),
                               act="relu" if is_bn else "none", tc=T if s == 1 else 0, bn=bn_l)
                if not is_bn:
                    ops.instnorm(t1, cout, t1, cout, stats, B, ho * wo, cout, relu_inner=True)
                sc = y
                if s != 1:
                    sc = buf(f"l{li}b{bi}.sc", ho, wo, cout)
                    if ys is not None:   # 1x1 stride 2 = 1x1 on block (0,0) of the space-to-depth tensor
                        ops.conv2d(w[p + ".down"], ys, 4 * cin, cin, B, ho, wo, sc, cout, tc=T, bn=bn_l)
                    else:
                        ops.conv2d(w[p + ".down"], y, cin, cin, B, h, wd, sc, cout, stride=s, pad=(0, 0))
                    if not is_bn:
                        ops.instnorm(sc, cout, sc, cout, stats, B, ho * wo, cout, relu_inner=False)
                out = buf(f"l{li}b{bi}.out", ho, wo, cout)
                if is_bn:
                    ops.conv2d(w[p + ".conv2"], t1, cout, cout, B, ho, wo, out, cout, pad=(1, 1), act="relu",
                               epi="add_relu", aux0=sc, aux0_cs=cout, tc=T, bn=bn_l)
                else:
                    t2 = buf(f"l{li}b{bi}.t2", ho, wo, cout)
                    ops.conv2d(w[p + ".conv2"], t1, cout, cout, B, ho, wo, t2, cout, pad=(1, 1), tc=T, bn=bn_l)
                    ops.instnorm(t2, cout, out, cout, stats, B, ho * wo, cout, relu_inner=True, relu_outer=True,
                                 res=sc, res_cs=cout)
                y, cin, h, wd = out, cout, ho, wo
        return y, h, wd

    # ---- CUDA-graph replay of the whole RAFT pass ----
    USE_GRAPH = _tc_mode("RAFT_GRAPH", 1)
    PARALLEL = _tc_mode("RAFT_PARALLEL", 1)    # two-stream schedule inside the RAFT pass
    _side = None

    def forward_packed_static(self, x2, hp, wp, iters=12):
        """forward_packed for a STATIC input buffer (update_cache's workspace): the ~200 launches of one RAFT pass
        are captured once into a CUDA graph (programmatic-dependent-launch edges included) and replayed per frame,
        which takes the Python/ctypes enqueue cost (measured 5.4 ms per 1080p frame, close to the GPU time) off
        the critical path.  Call 1 runs eagerly (allocations, weight packing), call 2 captures, later calls replay."""
        if not self.USE_GRAPH or torch.cuda.is_current_stream_capturing():
            # (under an outer capture - FrameEngine.infer_graphed - the launches go into that graph directly)
            return self.forward_packed(x2, hp, wp, iters)
        from . import _lib
        _dev_check(x2)
        self._prepare(x2.device)
        key = (x2.data_ptr(), hp, wp, iters, self._wkey)
        ent = self._graphs.get(key) if hasattr(self, "_graphs") else None
        if ent is None:
            self._graphs = {key: {}}          # one live configuration (a new size / new weights drop the old graph)
            return self.forward_packed(x2, hp, wp, iters)
        if "graph" not in ent:
            g = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            # thread_local: host threads of the caller (decoders, writers) may make CUDA calls while this thread captures
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                out = self.forward_packed(x2, hp, wp, iters)
            ent.update(graph=g, out=out, last=self.last, n=_lib.launch_count() - n0)
        ent["graph"].replay()
        _lib.lib().zt_add_launches(ent["n"])
        self.last = ent["last"]
        return ent["out"]

    # ---- full forward on packed inputs ----
    def forward_packed(self, x2, hp, wp, iters=12):
        """x2: [2][hp][wp][4] fp32 NHWC (image1, image2 normalised to [-1,1], padded).
        Returns flow_up NCHW [1,2,hp,wp] (a workspace tensor, valid until the next call)."""
        _dev_check(x2)
        self._prepare(x2.device)
        w, ws, T, TH = self.w, self.ws, self.TC_ENC, self.TC_HEAD
        assert hp % 8 == 0 and wp % 8 == 0
        h8, w8 = hp // 8, wp // 8
        N = h8 * w8
        # two-stream schedule (also captured as parallel graph branches): the context network + the per-frame GRU
        # planes run beside the feature network / correlation pyramid, and inside an iteration convf1/convf2 run
        # beside lookup/convc1/convc2.  Most RAFT launches fill 30..120 of the 148 SMs, so branches overlap.
        cur = torch.cuda.current_stream(x2.device)
        par = bool(self.PARALLEL)
        if par and (self._side is None or self._side.device != x2.device):
            self._side = torch.cuda.Stream(x2.device)
        side = self._side if par else cur

        def fork():
            if par:
                ev = torch.cuda.Event()
                ev.record(cur)
                side.wait_event(ev)

        def join():
            if par:
                ev = torch.cuda.Event()
                ev.record(side)
                cur.wait_event(ev)
        fused = bool(self.UPD_FUSED) and self.TC_UPD in (3, 4) and self.TC_HEAD == self.TC_UPD
        hx = ws.get("hx", (N, 384))
        # fused update kernel: the hidden state is its own dense tensor (fp32 master; the kernel keeps bf16 hi/lo copies)
        hbuf, hcs = (ws.get("h128", (N, 128)), 128) if fused else (hx, 384)
        gru_pre = {}
        fork()
        with torch.cuda.stream(side):
            c, ch, cwid = self._encoder("cnet", x2, 1, hp, wp, True)
            ops.conv2d(w["cnet.conv2.net"], c, 128, 128, 1, h8, w8, hbuf, hcs, act="tanh", tc=T, bn=64)
            ops.conv2d(w["cnet.conv2.inp"], c, 128, 128, 1, h8, w8, hx, 384, out_off=128, act="relu", tc=T, bn=64)
            if self.TC_UPD:
                for s_, pad_ in (("1", (0, 2)), ("2", (2, 0))):
                    for kind, co in (("zr", 256), ("q", 128)):
                        P = ws.get(f"gru.pre.{kind}{s_}", (N, co))
                        ops.conv2d(w[f"gru.{kind}{s_}.inp"], hx, 384, 128, 1, h8, w8, P, co, in0_off=128, pad=pad_,
                                   tc=self.TC_UPD, bn=64)
                        gru_pre[kind + s_] = P
        # feature network on both images
        f, fh, fw = self._encoder("fnet", x2, 2, hp, wp, False)
        assert (fh, fw) == (h8, w8)
        fmap = ws.get("fmap", (2 * N, 256))
        ops.conv2d(w["fnet.conv2"], f, 128, 128, 2, h8, w8, fmap, 256, tc=T, bn=64)
        # all-pairs correlation volume + pyramid (corr.py:13-27,52-60)
        hs, wsz = [h8], [w8]
        for _ in range(3):
            hs.append(hs[-1] // 2)
            wsz.append(wsz[-1] // 2)
        if hs[3] < 2 or wsz[3] < 2:
            raise RuntimeError("RAFT grid %dx%d is too small: the reference's 4-level lookup yields NaN when the "
                               "coarsest level is below 2x2 (needs a feature grid >= 16x16)" % (h8, w8))
        corr_dt = torch.bfloat16 if (self.CORR_BF16 and self.TC_CORR and self.CORR_FUSED) else torch.float32
        lv = [ws.get(f"corr{i}", (N, hs[i] * wsz[i]), corr_dt) for i in range(4)]
        if self.TC_CORR and self.CORR_FUSED:
            # volume + all three pooled levels from one tcgen05 accumulator (corr_gemm.cu)
            scratch = ws.get("corr_scratch", (ops.corr_pyramid_scratch_bytes(h8, w8) // 4,))
            ops.corr_pyramid(fmap, 256, h8, w8, lv, scratch)
        elif self.TC_CORR:
            # fmap1 x fmap2^T / sqrt(256) on the tensor cores (hi/lo split): fmap2 is packed as the
            # "weights" of a 1x1 convolution with cout = N
            cw = ws.bufs.get(("corr_w", N))
            if cw is None:
                cw = ops.RowsAsWeights(N, 256, 128, x2.device)
                ws.bufs[("corr_w", N)] = cw
            cw.pack(fmap[N:], 256)
            ops.conv2d(cw, fmap, 256, 256, 1, h8, w8, lv[0], N, scale=0.0625, tc=self.TC_CORR, bn=128)
        else:
            N4 = (N + 3) // 4 * 4
            f2t = ws.get("f2t", (256, N4))
            ops.transpose(fmap[N:], N, 256, 256, f2t, N4)
            cw = ConvW.__new__(ConvW)
            cw.w, cw.b, cw.kh, cw.kw, cw.cin, cw.cout, cw.w_cs = f2t, None, 1, 1, 256, N, N4
            ops.conv2d(cw, fmap, 256, 256, 1, h8, w8, lv[0], N, scale=0.0625)
        if not (self.TC_CORR and self.CORR_FUSED):
            for i in range(3):
                ops.avgpool2(lv[i], lv[i + 1], N, hs[i], wsz[i])
        # iteration state
        grid = ws.bufs.get(("grid", h8, w8))
        if grid is None:
            ys, xs = torch.meshgrid(torch.arange(h8, device=x2.device), torch.arange(w8, device=x2.device),
                                    indexing="ij")
            grid = torch.stack([xs, ys], -1).float().reshape(N, 2).contiguous()
            ws.bufs[("grid", h8, w8)] = grid
        coords1 = ws.get("coords1", (N, 2))
        flow4 = ws.get("flow4", (N, 4), zero=True)
        if fused:
            return self._finish_fused(w, ws, lv, hs, wsz, hbuf, gru_pre, coords1, flow4, fmap, h8, w8, hp, wp, iters, join)
        coords1.copy_(grid)
        corr = ws.get("corrfeat", (N, 324))
        c1 = ws.get("c1", (N, 256))
        c2f = ws.get("c2f", (N, 256))
        f1 = ws.get("f1", (N, 128))
        z = ws.get("z", (N, 128))
        rh = ws.get("rh", (N, 128))
        fhid = ws.get("fhid", (N, 256))
        delta = ws.get("delta", (N, 4), zero=True)
        ops.flow_update(coords1, None, 0, flow4, 4, hx, 384, 382, h8, w8)
        T = self.TC_UPD
        join()   # context network, hidden state and GRU planes are ready
        for it in range(iters):
            fork()
            with torch.cuda.stream(side):     # flow branch of the motion encoder (update.py:92-93)
                if T:
                    ops.raft_convf1(flow4, 4, w["enc.convf1.raw"][0], w["enc.convf1.raw"][1], f1, 128, h8, w8)
                else:
                    ops.conv2d(w["enc.convf1"], flow4, 4, 4, 1, h8, w8, f1, 128, pad=(3, 3), act="relu", tc=T, bn=64)
                ops.conv2d(w["enc.convf2"], f1, 128, 128, 1, h8, w8, c2f, 256, out_off=192, pad=(1, 1), act="relu", tc=T,
                           bn=32)
            ops.corr_lookup(lv, hs, wsz, coords1, corr, 324, N)   # correlation branch (update.py:90-91)
            ops.conv2d(w["enc.convc1"], corr, 324, 324, 1, h8, w8, c1, 256, act="relu", tc=T, bn=self.BN_C1)
            ops.conv2d(w["enc.convc2"], c1, 256, 256, 1, h8, w8, c2f, 256, pad=(1, 1), act="relu", tc=T, bn=self.BN_C2)
            join()
            ops.conv2d(w["enc.conv"], c2f, 256, 256, 1, h8, w8, hx, 384, out_off=256, pad=(1, 1), act="relu", tc=T,
                       bn=self.BN_ENC)
            for s, pad in (("1", (0, 2)), ("2", (2, 0))):
                if T:
                    ops.conv2d(w[f"gru.zr{s}.hm"], hx, 384, 128, 1, h8, w8, rh, 128, in1=hx, in1_cs=384, c1=128,
                               in1_off=256, pad=pad, epi="gru_zr", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128, tc=T,
                               bn=self.BN_ZR, bplane=gru_pre["zr" + s], bplane_cs=256)
                    ops.conv2d(w[f"gru.q{s}.hm"], rh, 128, 128, 1, h8, w8, hx, 384, in1=hx, in1_cs=384, c1=128,
                               in1_off=256, pad=pad, epi="gru_q", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128, tc=T,
                               bn=self.BN_Q, bplane=gru_pre["q" + s], bplane_cs=128)
                    continue
                ops.conv2d(w["gru.zr" + s], hx, 384, 384, 1, h8, w8, rh, 128, pad=pad, epi="gru_zr",
                           aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128, tc=T, bn=64)
                ops.conv2d(w["gru.q" + s], rh, 128, 128, 1, h8, w8, hx, 384, in1=hx, in1_cs=384, c1=256,
                           in1_off=128, pad=pad, epi="gru_q", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128, tc=T, bn=32)
            ops.conv2d(w["fh1"], hx, 384, 128, 1, h8, w8, fhid, 256, pad=(1, 1), act="relu", tc=TH, bn=self.BN_FH)
            if TH:     # flow head conv2 + coords1 += delta + flow outputs in one launch
                ops.flow_head2_update(fhid, 256, 256, w["fh2.raw"][0], w["fh2.raw"][1], coords1, flow4, 4, hx, 384, 382,
                                      h8, w8)
            else:
                ops.conv2d(w["fh2"], fhid, 256, 256, 1, h8, w8, delta, 4, pad=(1, 1), tc=TH)
                ops.flow_update(coords1, delta, 4, flow4, 4, hx, 384, 382, h8, w8)
        # convex upsampling mask only for the returned (last) prediction (raft.py:122-130, Q15)
        mh = ws.get("maskh", (N, 256))
        mask = ws.get("mask", (N, 576))
        ops.conv2d(w["mask0"], hx, 384, 128, 1, h8, w8, mh, 256, pad=(1, 1), act="relu", tc=TH, bn=64)
        ops.conv2d(w["mask2"], mh, 256, 256, 1, h8, w8, mask, 576, scale=0.25, tc=TH, bn=64)
        flow_up = ws.get("flow_up", (1, 2, hp, wp))
        ops.convex_upsample(flow4, 4, mask, 576, flow_up, h8, w8)
        self.last = dict(fmap=fmap, levels=lv, hs=hs, ws=wsz, flow4=flow4, hx=hx, corr=corr, N=N, h8=h8, w8=w8)
        return flow_up

    def _finish_fused(self, w, ws, lv, hs, wsz, h, gru_pre, coords1, flow4, fmap, h8, w8, hp, wp, iters, join):
        """The update loop (raft.py:111-120) as ONE persistent kernel, then mask head + convex upsampling of the last
        prediction.  Activations between the layers live as bf16 hi/lo pairs (TMA-fed tensor cores)."""
        N, bf = h8 * w8, torch.bfloat16
        bufs = dict(h=h, coords1=coords1, flow4=flow4, z=ws.get("z", (N, 128)), fhid=ws.get("fhid", (N, 256)),
                    counter=ws.get("ru.counter", (4,), torch.int32, zero=True))
        for name, cn in (("corr", 384), ("c1", 256), ("f1", 128), ("c2f", 256), ("hm", 256), ("rh", 128)):
            for sfx in ("_hi", "_lo"):
                # corr: channels 324..383 are never written and must read as zero (K padding of convc1)
                bufs[name + sfx] = ws.get("ru." + name + sfx, (N, cn), bf, zero=True)
        wl = [w["enc.convc1"], w["enc.convf2"], w["enc.convc2"], w["enc.conv"], w["gru.zr1.hm"], w["gru.q1.hm"],
              w["gru.zr2.hm"], w["gru.q2.hm"], w["fh1"]]
        join()   # context network, hidden state and GRU planes are ready
        ops.raft_update(h8, w8, iters, bufs, (gru_pre["zr1"], gru_pre["q1"], gru_pre["zr2"], gru_pre["q2"]), lv, hs, wsz,
                        wl, self.FUSED_BN, w["enc.convf1.raw"][0], w["enc.convf1.raw"][1], w["fh2.raw"][0], w["fh2.raw"][1],
                        max_ctas=self.FUSED_MAX_CTAS, fp16=(self.TC_UPD == 4))
        TH = self.TC_HEAD
        mh = ws.get("maskh", (N, 256))
        mask = ws.get("mask", (N, 576))
        ops.conv2d(w["mask0"], h, 128, 128, 1, h8, w8, mh, 256, pad=(1, 1), act="relu", tc=TH, bn=64)
        ops.conv2d(w["mask2"], mh, 256, 256, 1, h8, w8, mask, 576, scale=0.25, tc=TH, bn=64)
        flow_up = ws.get("flow_up", (1, 2, hp, wp))
        ops.convex_upsample(flow4, 4, mask, 576, flow_up, h8, w8)
        self.last = dict(fmap=fmap, levels=lv, hs=hs, ws=wsz, flow4=flow4, hx=None, corr=None, N=N, h8=h8, w8=w8, h=h)
        return flow_up

    def forward(self, image1, image2, iters=12):
        """API-compatible RAFT.forward on NCHW [1,3,H,W] images in 0..255 (raft.py:77-130).
        Input padding / normalisation / packing use torch device ops here; the per-frame hot
        path (update_cache) uses the fused kernels instead."""
        _dev_check(image1)
        import torch.nn.functional as F
        assert image1.shape[0] == 1, "batch 1 only (the reference CLIs use batch 1)"
        ht, wd = image1.shape[-2:]
        ph, pw = (((ht // 8) + 1) * 8 - ht) % 8, (((wd // 8) + 1) * 8 - wd) % 8
        pad = [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2]
        x = torch.cat([image1, image2], 0).float()
        x = F.pad(x, pad, mode="replicate")
        x = 2 * (x / 255.0) - 1.0
        x = F.pad(x.permute(0, 2, 3, 1), [0, 1]).contiguous()
        hp, wp = ht + ph, wd + pw
        flow_up = self.forward_packed(x, hp, wp, iters).clone()
        flow_low = self.last["flow4"][:, :2].reshape(1, self.last["h8"], self.last["w8"], 2).permute(0, 3, 1, 2)
        return flow_low.contiguous(), flow_up


# ----------------------------------------------------------------------------------------
# Retinex nets + update_cache
# ----------------------------------------------------------------------------------------
class FrameEngine:
    """Finetunemodel.forward / update_cache (model/model.py:312-384) on the kernels."""

    EPS = 1e-4
    # plain bf16 tensor-core products for the three Retinex nets: measured 93 dB PSNR / 2e-4 px
    # EPE against the fp32 path over a 12-frame recurrence (profiles/r01_precision_sweep.txt),
    # far inside the 50 dB / 0.01 px gates.  RAFT itself needs the hi/lo split (0.07 px in bf16).
    # mode 2 (default): bf16 NHWC activations between layers, TMA-fed tcgen05 kernel (csrc/conv_tma.cu)
    TC_D1 = _tc_mode("D1", 2)
    TC_ENH = _tc_mode("ENH", 2)
    TC_D2 = _tc_mode("D2", 2)
    HEAD_FUSED = int(os.environ.get("ZT_HEAD_FUSED", "1"))   # denoisers' 1x1 heads in the 3x3 layer's epilogue (conv_tma OUT = 3)

    def __init__(self, model):
        self.model = model
        self._wkey = None
        self.ws = None

    def _prepare(self, device):
        m = self.model
        key = (str(device), _param_version(m.enhance), _param_version(m.denoise_1), _param_version(m.denoise_2))
        if key == self._wkey:
            return
        w = {}
        for n, d, cin_pad in (("d1", m.denoise_1, 4), ("d2", m.denoise_2, 12)):
            w[n + ".c1"] = ConvW(d.conv1.weight.to(device), d.conv1.bias.to(device), cin_pad=cin_pad)
            w[n + ".c2"] = ConvW(d.conv2.weight.to(device), d.conv2.bias.to(device))
            w[n + ".c3"] = ConvW(d.conv3.weight.to(device), d.conv3.bias.to(device))
        e = m.enhance
        w["e.in"] = ConvW(e.in_conv[0].weight.to(device), e.in_conv[0].bias.to(device), cin_pad=12)
        bnm = e.conv[1]
        w["e.blk"] = ConvW(e.conv[0].weight.to(device), e.conv[0].bias.to(device),
                           bn=(bnm.weight.to(device), bnm.bias.to(device), bnm.running_mean.to(device),
                               bnm.running_var.to(device)), eps=bnm.eps)
        w["e.out"] = ConvW(e.out_conv[0].weight.to(device), e.out_conv[0].bias.to(device))
        self.w, self._wkey = w, key
        if self.ws is None or self.ws.device != device:
            self.ws = _Workspace(device)

    # ---- nets ----
    def denoise1_L2(self, x, H, W, tag=""):
        """L2 = clamp((x+eps) - denoise_1(x+eps), eps, 1)   (model.py:313-316).  `tag` names a private set of
        activation buffers (training keeps them for the backward pass)."""
        w, ws, HW = self.w, self.ws, H * W
        d = ws.get("d4" + tag, (HW, 4))
        if self.TC_D1 == 2:
            bf = torch.bfloat16
            xp, a1, a2 = ws.get("xp8", (HW, 8), bf), ws.get("b48a", (HW, 48), bf), ws.get("b48b", (HW, 48), bf)
            ops.pack_nhwc_bf16([x], 3, self.EPS, xp, 8, HW)
            ops.conv2d_bf16(w["d1.c1"], xp, 8, H, W, a1, 48, act="lrelu")
            if self.HEAD_FUSED:    # conv3 (1x1, 48 -> 3) applied in conv2's epilogue: a2 is never written
                ops.conv2d_bf16(w["d1.c2"], a1, 48, H, W, d, 4, act="lrelu", head=w["d1.c3"])
            else:
                ops.conv2d_bf16(w["d1.c2"], a1, 48, H, W, a2, 48, act="lrelu")
                ops.conv2d_bf16(w["d1.c3"], a2, 48, H, W, d, 4, out_f32=True)
            L2 = torch.empty_like(x)
            ops.retinex_l2(x, d, 4, L2, HW, self.EPS)
            return L2
        xp = ws.get("xp" + tag, (HW, 4))
        a1 = ws.get("a48a" + tag, (HW, 48))
        a2 = ws.get("a48b" + tag, (HW, 48))
        ops.pack_nhwc([x], 3, self.EPS, xp, 4, HW)
        ops.conv2d(w["d1.c1"], xp, 4, 4, 1, H, W, a1, 48, pad=(1, 1), act="lrelu", tc=self.TC_D1)
        ops.conv2d(w["d1.c2"], a1, 48, 48, 1, H, W, a2, 48, pad=(1, 1), act="lrelu", tc=self.TC_D1)
        ops.conv2d(w["d1.c3"], a2, 48, 48, 1, H, W, d, 4, tc=self.TC_D1)
        L2 = torch.empty_like(x)
        ops.retinex_l2(x, d, 4, L2, HW, self.EPS)
        return L2

    def enhance_H2(self, x, H3_wp, s3_wp, L2, H, W):
        """s2 = enhance(cat[H3_wp, s3_wp, L2]); H2 = clamp((x+eps)/s2, eps, 1)  (model.py:326-328)."""
        w, ws, HW = self.w, self.ws, H * W
        s4 = ws.get("s4", (HW, 4))
        if self.TC_ENH == 2:
            bf = torch.bfloat16
            ep, cur, nxt = ws.get("ep16", (HW, 16), bf), ws.get("b64a", (HW, 64), bf), ws.get("b64b", (HW, 64), bf)
            ops.pack_nhwc_bf16([H3_wp, s3_wp, L2], 3, 0.0, ep, 16, HW)
            ops.conv2d_bf16(w["e.in"], ep, 16, H, W, cur, 64, act="relu")
            for _ in range(3):  # one shared conv+BN applied three times (model.py:60-67, Q5)
                ops.conv2d_bf16(w["e.blk"], cur, 64, H, W, nxt, 64, act="relu", res=cur, res_cs=64)
                cur, nxt = nxt, cur
            ops.conv2d_bf16(w["e.out"], cur, 64, H, W, s4, 4, act="sigmoid", out_f32=True)
            s2 = torch.empty_like(x)
            H2 = torch.empty_like(x)
            ops.retinex_h2(x, s4, 4, s2, H2, HW, self.EPS)
            return s2, H2
        ep = ws.get("ep12", (HW, 12))
        fa = ws.get("f64a", (HW, 64))
        fb = ws.get("f64b", (HW, 64))
        ops.pack_nhwc([H3_wp, s3_wp, L2], 3, 0.0, ep, 12, HW)
        ops.conv2d(w["e.in"], ep, 12, 12, 1, H, W, fa, 64, pad=(1, 1), act="relu", tc=self.TC_ENH)
        cur, nxt = fa, fb
        for _ in range(3):  # one shared conv+BN applied three times (model.py:60-67, Q5)
            ops.conv2d(w["e.blk"], cur, 64, 64, 1, H, W, nxt, 64, pad=(1, 1), act="relu", epi="add",
                       aux0=cur, aux0_cs=64, tc=self.TC_ENH)
            cur, nxt = nxt, cur
        ops.conv2d(w["e.out"], cur, 64, 64, 1, H, W, s4, 4, pad=(1, 1), act="sigmoid", tc=self.TC_ENH)
        s2 = torch.empty_like(x)
        H2 = torch.empty_like(x)
        ops.retinex_h2(x, s4, 4, s2, H2, HW, self.EPS)
        return s2, H2

    def denoise2_H3(self, H3_wp, s3_wp, H2, s2, H, W, tag="", out=None):
        """H5 = clamp(cat[H2,s2] - denoise_2(cat[H3_wp,s3_wp,H2,s2]), eps, 1)  (model.py:334-337).
        out = (H3, s3) writes into caller-owned buffers (the static state of infer_graphed)."""
        w, ws, HW = self.w, self.ws, H * W
        d = ws.get("d8" + tag, (HW, 8))
        if self.TC_D2 == 2:
            bf = torch.bfloat16
            dp, a1, a2 = ws.get("dp16", (HW, 16), bf), ws.get("b48a", (HW, 48), bf), ws.get("b48b", (HW, 48), bf)
            ops.pack_nhwc_bf16([H3_wp, s3_wp, H2, s2], 3, 0.0, dp, 16, HW)
            ops.conv2d_bf16(w["d2.c1"], dp, 16, H, W, a1, 48, act="lrelu")
            if self.HEAD_FUSED:
                ops.conv2d_bf16(w["d2.c2"], a1, 48, H, W, d, 8, act="lrelu", head=w["d2.c3"])
            else:
                ops.conv2d_bf16(w["d2.c2"], a1, 48, H, W, a2, 48, act="lrelu")
                ops.conv2d_bf16(w["d2.c3"], a2, 48, H, W, d, 8, out_f32=True)
            H3, s3 = out if out is not None else (torch.empty_like(H2), torch.empty_like(H2))
            ops.retinex_h5(H2, s2, d, 8, H3, s3, HW, self.EPS)
            return H3, s3
        dp = ws.get("dp12" + tag, (HW, 12))
        a1 = ws.get("a48a" + tag, (HW, 48))
        a2 = ws.get("a48b" + tag, (HW, 48))
        ops.pack_nhwc([H3_wp, s3_wp, H2, s2], 3, 0.0, dp, 12, HW)
        ops.conv2d(w["d2.c1"], dp, 12, 12, 1, H, W, a1, 48, pad=(1, 1), act="lrelu", tc=self.TC_D2)
        ops.conv2d(w["d2.c2"], a1, 48, 48, 1, H, W, a2, 48, pad=(1, 1), act="lrelu", tc=self.TC_D2)
        ops.conv2d(w["d2.c3"], a2, 48, 48, 1, H, W, d, 8, tc=self.TC_D2)
        H3, s3 = out if out is not None else (torch.empty_like(H2), torch.empty_like(H2))
        ops.retinex_h5(H2, s2, d, 8, H3, s3, HW, self.EPS)
        return H3, s3

    # ---- update_cache (model.py:346-384) ----
    def raft_inputs(self, last_H3, L2, of_scale):
        """Resize, *255, quantise+equalise image2, normalise, pad, pack -> [2][hp][wp][4]."""
        ws = self.ws
        H, W = last_H3.shape[-2:]
        ht, wd = H // of_scale, W // of_scale
        hp, wp = (ht + 7) // 8 * 8, (wd + 7) // 8 * 8
        x2 = ws.get("raft_in", (2, hp, wp, 4))
        q = ws.get("l2_u8", (3, ht, wd), torch.uint8)
        hist = ws.get("hist", (768,), torch.int32)
        lut = ws.get("lut", (768,), torch.uint8)
        ops.resize255_image1(last_H3, H, W, ht, wd, x2[0], hp, wp)
        ops.resize255_quant_hist(L2, H, W, ht, wd, q, hist)
        ops.equalize_lut(hist, lut)
        ops.equalize_apply_pack(q, lut, ht, wd, hp, wp, x2[1])
        return x2, hp, wp

    def update_cache(self, last_H3, last_s3, L2, of_scale):
        for t in (last_H3, last_s3, L2):
            _dev_check(t)
        self._prepare(L2.device)
        assert last_H3.shape[0] == 1 and last_H3.shape[1] == 3, "batch 1, 3 channels"
        last_H3, last_s3, L2 = last_H3.contiguous(), last_s3.contiguous(), L2.contiguous()
        H, W = last_H3.shape[-2:]
        x2, hp, wp = self.raft_inputs(last_H3, L2, of_scale)
        flow_up = self.model.raft.engine().forward_packed_static(x2, hp, wp, 12)
        H3_wp = torch.empty_like(last_H3)
        s3_wp = torch.empty_like(last_s3)
        ops.warp_pair(flow_up, hp, wp, last_H3, last_s3, H3_wp, s3_wp, 3, H, W)
        return H3_wp, s3_wp

    # ---- Finetunemodel.forward (model.py:312-340) ----
    def infer(self, x, is_new_seq, last_H3, last_s3, of_scale):
        _dev_check(x)
        assert x.dim() == 4 and x.shape[0] == 1 and x.shape[1] == 3, "expects [1,3,H,W] (the CLIs use batch 1)"
        x = x.contiguous().float()
        self._prepare(x.device)
        H, W = x.shape[-2:]
        L2 = self.denoise1_L2(x, H, W)
        if is_new_seq:
            zeros = self.ws.get("zeros3", (1, 3, H, W), zero=True)
            H3_wp = s3_wp = zeros
        else:
            H3_wp, s3_wp = self.update_cache(last_H3, last_s3, L2, of_scale)
        s2, H2 = self.enhance_H2(x, H3_wp, s3_wp, L2, H, W)
        if is_new_seq:  # inference-only substitution (model.py:330-332, Q12)
            H3_wp = s3_wp = H2
        H3, s3 = self.denoise2_H3(H3_wp, s3_wp, H2, s2, H, W)
        return dict(L2=L2, s2=s2, H2=H2, H3=H3, s3=s3, H3_wp=H3_wp, s3_wp=s3_wp)

    # ---- steady-state frame as ONE CUDA graph (streaming pipelines) ----
    USE_FRAME_GRAPH = int(os.environ.get("ZT_FRAME_GRAPH", "1"))

    def _steady_body(self, x, st, of_scale):
        """Finetunemodel.forward for is_new_seq == False (model.py:312-340) reading the recurrent state from, and
        writing the new state into, the static buffers st["H3"], st["s3"] (the state is read early - resize, warp -
        and written by the very last kernel of the frame, so one buffer pair serves both)."""
        H, W = x.shape[-2:]
        L2 = self.denoise1_L2(x, H, W)
        H3_wp, s3_wp = self.update_cache(st["H3"], st["s3"], L2, of_scale)
        s2, H2 = self.enhance_H2(x, H3_wp, s3_wp, L2, H, W)
        H3, s3 = self.denoise2_H3(H3_wp, s3_wp, H2, s2, H, W, out=(st["H3"], st["s3"]))
        return dict(L2=L2, s2=s2, H2=H2, H3=H3, s3=s3, H3_wp=H3_wp, s3_wp=s3_wp)

    def infer_graphed(self, x, last_H3, last_s3, of_scale):
        """Steady-state frame for a STATIC input buffer `x` (same address every call): the ~250 launches of the
        frame - Retinex nets, RAFT pass with its two-stream branches, warp - are captured once into a CUDA graph and
        replayed, which takes the per-frame Python / ctypes enqueue cost (~2.3 ms at 1080p, comparable with the GPU
        time of the frame) down to one graph launch.  Call 1 runs eagerly (allocations, weight packing), call 2
        captures, later calls replay.  The returned tensors are static: valid until the next call."""
        from . import _lib
        _dev_check(x)
        assert x.dim() == 4 and x.shape[0] == 1 and x.shape[1] == 3 and x.dtype == torch.float32 and x.is_contiguous()
        self._prepare(x.device)
        reng = self.model.raft.engine()
        reng._prepare(x.device)
        key = (x.data_ptr(), tuple(x.shape), of_scale, self._wkey, reng._wkey)
        st = getattr(self, "_fg", None)
        if st is None or st["key"] != key:
            st = self._fg = dict(key=key, H3=torch.empty_like(x), s3=torch.empty_like(x), n=0)
        if last_H3.data_ptr() != st["H3"].data_ptr():      # state produced elsewhere (new-sequence frame, eager call)
            st["H3"].copy_(last_H3)
            st["s3"].copy_(last_s3)
        if not self.USE_FRAME_GRAPH or st["n"] == 0:
            st["n"] += 1
            return self._steady_body(x, st, of_scale)
        if "graph" not in st:
            g = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                out = self._steady_body(x, st, of_scale)
            st.update(graph=g, out=out, launches=_lib.launch_count() - n0)
        st["graph"].replay()
        _lib.lib().zt_add_launches(st["launches"])
        return st["out"]


# ----------------------------------------------------------------------------------------
# Training step: Network.forward / LossFunction.forward and their backward pass
# ----------------------------------------------------------------------------------------
SMOOTH_OFFSETS = ((1, 0), (0, 1), (1, 1), (1, -1), (2, 0), (0, 2), (2, 1), (2, -1), (1, 2), (1, -2), (2, 2), (2, -2))
LOSS_NAMES = ("enh_s2 enh_norm smooth tv res1_a res1_b res1_c res1_d res2_a res2_b res2_c res2_d "
              "color ill inter1 inter2 var").split()
# trainable parameters in the order the gradients are returned (== Network.named_parameters() minus raft.*)
TRAIN_PARAMS = ("enhance.in_conv.0.weight enhance.in_conv.0.bias enhance.conv.0.weight enhance.conv.0.bias "
                "enhance.conv.1.weight enhance.conv.1.bias enhance.out_conv.0.weight enhance.out_conv.0.bias "
                "denoise_1.conv1.weight denoise_1.conv1.bias denoise_1.conv2.weight denoise_1.conv2.bias "
                "denoise_1.conv3.weight denoise_1.conv3.bias denoise_2.conv1.weight denoise_2.conv1.bias "
                "denoise_2.conv2.weight denoise_2.conv2.bias denoise_2.conv3.weight denoise_2.conv3.bias").split()


class TrainEngine(FrameEngine):
    """Network.forward (model/model.py:144-203), LossFunction.forward (loss.py:23-78) and the backward pass
    of train.py:126-131 on the kernels.  The nets run the bf16 hi/lo split here (fp32-grade products): the
    loss gate is 1e-3 relative, the gradient gate 1e-3 relative L2 per tensor."""

    TC_D1 = _tc_mode("TRAIN_D1", 3)
    TC_ENH = _tc_mode("TRAIN_ENH", 3)
    TC_D2 = _tc_mode("TRAIN_D2", 3)
    TC_BWD = _tc_mode("TRAIN_BWD", 3)

    # weights of the tensor-core paths are re-packed by a device kernel after every optimiser step (ops.DevConvW)
    DEV_PACK = int(os.environ.get("ZT_TRAIN_DEVPACK", "1"))

    def _dev_pack_ok(self, device):
        m = self.model
        return (self.DEV_PACK and all(t in (3, 4) for t in (self.TC_D1, self.TC_ENH, self.TC_D2, self.TC_BWD)) and
                all(p.device == device and p.dtype == torch.float32 and p.is_contiguous()
                    for mod in (m.enhance, m.denoise_1, m.denoise_2) for p in mod.parameters()))

    def _prepare(self, device):
        m = self.model
        key = (str(device), _param_version(m.enhance), _param_version(m.denoise_1), _param_version(m.denoise_2))
        if self._dev_pack_ok(device):
            ptrs = tuple(p.data_ptr() for mod in (m.enhance, m.denoise_1, m.denoise_2) for p in mod.parameters())
            if getattr(self, "_devw_ptrs", None) != ptrs:        # first use (or the parameters moved): build the packers
                from .ops import DevConvW
                e, w, wb = m.enhance, {}, {}
                for n, d, cp in (("d1", m.denoise_1, 4), ("d2", m.denoise_2, 12)):
                    w[n + ".c1"] = DevConvW(d.conv1.weight, d.conv1.bias, cin_pad=cp)
                    w[n + ".c2"] = DevConvW(d.conv2.weight, d.conv2.bias)
                    w[n + ".c3"] = DevConvW(d.conv3.weight, d.conv3.bias)
                w["e.in"] = DevConvW(e.in_conv[0].weight, e.in_conv[0].bias, cin_pad=12)
                w["e.blk_raw"] = DevConvW(e.conv[0].weight, e.conv[0].bias)
                w["e.out"] = DevConvW(e.out_conv[0].weight, e.out_conv[0].bias)
                wb["e.out"] = DevConvW(e.out_conv[0].weight, None, cin_pad=4, dgrad=True)
                wb["e.blk"] = DevConvW(e.conv[0].weight, None, dgrad=True)
                wb["d1.c3"] = DevConvW(m.denoise_1.conv3.weight, None, cin_pad=4, dgrad=True)
                wb["d1.c2"] = DevConvW(m.denoise_1.conv2.weight, None, dgrad=True)
                wb["d2.c3"] = DevConvW(m.denoise_2.conv3.weight, None, cin_pad=8, dgrad=True)
                wb["d2.c2"] = DevConvW(m.denoise_2.conv2.weight, None, dgrad=True)
                wb["d2.c1"] = DevConvW(m.denoise_2.conv1.weight, None, dgrad=True)
                self.w, self._wbwd, self._devw_ptrs, self._wkey, self._wg = w, wb, ptrs, key, None
                self._eval_key = None
                if self.ws is None or self.ws.device != device:
                    self.ws = _Workspace(device)
            elif key != self._wkey:                              # an optimiser step happened: one small kernel per layer
                for d_ in (self.w, self._wbwd):
                    for wt in d_.values():
                        if getattr(wt, "dynamic", False):
                            wt.repack()
                self._wkey = key
            return
        self._devw_ptrs = None
        fresh = key != self._wkey
        super()._prepare(device)
        if fresh:   # train-mode BatchNorm needs the un-folded block convolution
            e = m.enhance
            self.w["e.blk_raw"] = ConvW(e.conv[0].weight.to(device), e.conv[0].bias.to(device))
            self._wbwd = None

    def _eval_block(self, device):
        """Eval-mode forward of the training model (train.py:138): the BatchNorm-folded block convolution is packed on
        the host (rare: once per evaluation pass)."""
        if getattr(self, "_devw_ptrs", None) is None:
            return
        if getattr(self, "_eval_key", None) != self._wkey:
            e = self.model.enhance
            bnm = e.conv[1]
            self.w["e.blk"] = ConvW(e.conv[0].weight.to(device), e.conv[0].bias.to(device),
                                    bn=(bnm.weight.to(device), bnm.bias.to(device), bnm.running_mean.to(device),
                                        bnm.running_var.to(device)), eps=bnm.eps)
            self._eval_key = self._wkey

    # L_pred = x - denoise_1(x) without clamp (model.py:149-150)
    def denoise1_residual(self, x, H, W, tag=""):
        w, ws, HW = self.w, self.ws, H * W
        xp, a1, a2, d = (ws.get("xp" + tag, (HW, 4)), ws.get("a48a" + tag, (HW, 48)), ws.get("a48b" + tag, (HW, 48)),
                         ws.get("d4" + tag, (HW, 4)))
        ops.pack_nhwc([x], 3, 0.0, xp, 4, HW)
        ops.conv2d(w["d1.c1"], xp, 4, 4, 1, H, W, a1, 48, pad=(1, 1), act="lrelu", tc=self.TC_D1)
        ops.conv2d(w["d1.c2"], a1, 48, 48, 1, H, W, a2, 48, pad=(1, 1), act="lrelu", tc=self.TC_D1)
        ops.conv2d(w["d1.c3"], a2, 48, 48, 1, H, W, d, 4, tc=self.TC_D1)
        out = torch.empty_like(x)
        ops.sub_clamp(x, 0.0, d, 4, out, 3, HW)
        return out

    def enhance_train(self, x, H3_wp, s3_wp, L2, H, W):
        """Enhancer.forward with BatchNorm in TRAIN mode (batch statistics, running stats updated three
        times - model.py:60-67,74-81 under train.py:116), then H2 = clamp((x+eps)/s2, eps, 1).  Every
        intermediate (pre-BN conv outputs, block outputs, batch statistics) is kept for the backward pass."""
        w, ws, HW = self.w, self.ws, H * W
        bn = self.model.enhance.conv[1]
        ep = ws.get("ep12", (HW, 12))
        cur = [ws.get("e.cur%d" % k, (HW, 64)) for k in range(4)]
        t = [ws.get("e.t%d" % k, (HW, 64)) for k in range(3)]
        st = [ws.get("e.stats%d" % k, (128,), torch.float64) for k in range(3)]
        s4 = ws.get("s4", (HW, 4))
        ops.pack_nhwc([H3_wp, s3_wp, L2], 3, 0.0, ep, 12, HW)
        ops.conv2d(w["e.in"], ep, 12, 12, 1, H, W, cur[0], 64, pad=(1, 1), act="relu", tc=self.TC_ENH)
        for k in range(3):
            ops.conv2d(w["e.blk_raw"], cur[k], 64, 64, 1, H, W, t[k], 64, pad=(1, 1), tc=self.TC_ENH)
            ops.bn_train(t[k], 64, cur[k + 1], 64, cur[k], 64, st[k], bn.weight, bn.bias, bn.running_mean,
                         bn.running_var, HW, 64, bn.eps, bn.momentum)
            bn.num_batches_tracked += 1
        bump_param_generation((bn.running_mean, bn.running_var))   # written through raw pointers
        ops.conv2d(w["e.out"], cur[3], 64, 64, 1, H, W, s4, 4, pad=(1, 1), act="sigmoid", tc=self.TC_ENH)
        s2, H2 = torch.empty_like(x), torch.empty_like(x)
        ops.retinex_h2(x, s4, 4, s2, H2, HW, self.EPS)
        return s2, H2

    def forward_train(self, x, is_new_seq, last_H3, last_s3, of_scale, training):
        _dev_check(x)
        assert x.dim() == 4 and x.shape[0] == 1 and x.shape[1] == 3, "expects [1,3,H,W] (train.py uses batch 1)"
        x = x.contiguous().float()
        self._prepare(x.device)
        H, W = x.shape[-2:]
        h, wd = H // 2, W // 2
        eps = self.EPS
        L11, L12 = ops.pair_downsample(x, eps)
        L_pred1 = self.denoise1_residual(L11, h, wd, ".1a")
        L_pred2 = self.denoise1_residual(L12, h, wd, ".1b")
        L2 = self.denoise1_L2(x, H, W, ".1f")
        if is_new_seq:
            H3_wp = s3_wp = torch.zeros_like(L2)
            H31_wp = H32_wp = s31_wp = s32_wp = torch.zeros_like(L11)
        else:
            H3_wp, s3_wp = self.update_cache(last_H3, last_s3, L2, of_scale)
            H31_wp, H32_wp = ops.pair_downsample(H3_wp)
            s31_wp, s32_wp = ops.pair_downsample(s3_wp)
        if training:
            s2, H2 = self.enhance_train(x, H3_wp, s3_wp, L2, H, W)
        else:
            self._eval_block(x.device)
            s2, H2 = self.enhance_H2(x, H3_wp, s3_wp, L2, H, W)
        s21, s22 = ops.pair_downsample(s2)
        H11 = ops.div_clamp(L11, 0.0, s21, eps, 1.0)
        H12 = ops.div_clamp(L12, 0.0, s22, eps, 1.0)
        H13, s13 = self.denoise2_H3(H31_wp, s31_wp, H11, s21, h, wd, ".a")
        H14, s14 = self.denoise2_H3(H32_wp, s32_wp, H12, s22, h, wd, ".b")
        H3, s3 = self.denoise2_H3(H3_wp, s3_wp, H2, s2, H, W, ".f")
        H3_pred, H4_pred = torch.cat([H13, s13], 1), torch.cat([H14, s14], 1)
        d12 = ops.texture_diff(L_pred1, L_pred2)
        H3d1, H3d2 = ops.pair_downsample(H3)
        dH3 = ops.texture_diff(H3d1, H3d2)
        H1 = ops.div_clamp(L2, 0.0, s2, 0.0, 1.0)
        H2_blur, H3_blur = ops.blur21(H1), ops.blur21(H3)
        self.last_wp = (H3_wp, s3_wp)
        self.ctx = dict(x=x, H=H, W=W, L11=L11, L12=L12, training=training)
        return (L_pred1, L_pred2, L2, s2, s21, s22, H2, H11, H12, H13, s13, H14, s14, H3, s3, H3_pred, H4_pred,
                d12, dH3, H2_blur, H3_blur, H3d1, H3d2)

    def loss_terms(self, x, outs, is_WB):
        """The 17 weighted terms of LossFunction.forward (loss.py:46-77) as a device fp64 vector."""
        (L_pred1, L_pred2, L2, s2, s21, s22, H2, H11, H12, H13, s13, H14, s14, H3, s3, H3_pred, H4_pred,
         d12, dH3, H2_blur, H3_blur, H3d1, H3d2) = outs
        x = x.contiguous().float()
        H, W = x.shape[-2:]
        h, wd = H // 2, W // 2
        ws = self.ws
        acc = torch.zeros(32, dtype=torch.float64, device=x.device)
        chansum = ws.get("chansum", (4,), torch.float64)
        ycc = ws.get("loss_ycc", (1, 3, H, W))
        avg1, avg2 = ws.get("loss_avg1", (1, 3, H, W)), ws.get("loss_avg2", (1, 3, H, W))
        ops.loss_illum(L2, s2, is_WB, chansum, acc, 0)
        ops.loss_smooth(L2, s2, ycc, acc, 2)
        ops.loss_tv(s2, acc, 14)
        L11, L12 = ops.pair_downsample(x, 1e-9)       # loss.py:24,51: input + 1e-9
        den1, den2 = ops.pair_downsample(L2)
        ops.mse_sum(L11, L_pred2, acc, 16)
        ops.mse_sum(L12, L_pred1, acc, 17)
        ops.mse_sum(L_pred1, den1, acc, 18)
        ops.mse_sum(L_pred2, den2, acc, 19)
        ops.mse_sum(H13, H12, acc, 20)
        ops.mse_sum(s13, s22, acc, 20)
        ops.mse_sum(H14, H11, acc, 21)
        ops.mse_sum(s14, s21, acc, 21)
        ops.mse_sum(H13, H3d1, acc, 22)
        ops.mse_sum(H14, H3d2, acc, 23)
        ops.mse_sum(H2_blur, H3_blur, acc, 24)
        ops.mse_sum(s2, s3, acc, 25)
        ops.loss_inter(H3d1, H3d2, dH3, acc, 26)
        ops.loss_var(H2, H3, avg1, avg2, acc, 28)
        n3, nh = 3.0 * H * W, 3.0 * h * wd
        sm = sum(2.0 * acc[2 + i] / ((H - dy) * (W - abs(dx))) for i, (dy, dx) in enumerate(SMOOTH_OFFSETS))
        tv = 2.0 * (acc[14] / ((H - 1) * W) + acc[15] / (H * (W - 1)))
        terms = [700 * acc[0] / n3, 1000 * acc[1] / n3, 5 * sm, 1600 * tv,
                 1000 * acc[16] / nh, 1000 * acc[17] / nh, 1000 * acc[18] / nh, 1000 * acc[19] / nh,
                 1000 * acc[20] / (2 * nh), 1000 * acc[21] / (2 * nh), 1000 * acc[22] / nh, 1000 * acc[23] / nh,
                 10000 * acc[24] / n3, 1000 * acc[25] / n3, 10000 * acc[26] / nh, 10000 * acc[27] / nh,
                 1000 * acc[28] / n3]
        if getattr(self, "ctx", None) is not None:
            self.ctx.update(outs=outs, is_WB=is_WB, L11e=L11, L12e=L12, den1=den1, den2=den2)
        return torch.stack(terms)

    # ---- backward (train.py:129 loss.backward()) ----
    def _prepare_bwd(self, device):
        if getattr(self, "_devw_ptrs", None) is not None:      # device-packed data-gradient weights exist already
            if getattr(self, "_wg", None) is None:
                self._make_wgrads(device)
            return
        if getattr(self, "_wbwd", None) is not None:
            return
        m, w = self.model, {}
        e, d1, d2 = m.enhance, m.denoise_1, m.denoise_2
        w["e.out"] = ops.dgrad_weights(e.out_conv[0].weight.to(device), cin_pad=4)
        w["e.blk"] = ops.dgrad_weights(e.conv[0].weight.to(device))
        w["d1.c3"] = ops.dgrad_weights(d1.conv3.weight.to(device), cin_pad=4)
        w["d1.c2"] = ops.dgrad_weights(d1.conv2.weight.to(device))
        w["d2.c3"] = ops.dgrad_weights(d2.conv3.weight.to(device), cin_pad=8)
        w["d2.c2"] = ops.dgrad_weights(d2.conv2.weight.to(device))
        w["d2.c1"] = ops.dgrad_weights(d2.conv1.weight.to(device))
        self._wbwd = w
        if getattr(self, "_wg", None) is None:
            self._make_wgrads(device)

    def _make_wgrads(self, device):
        m = self.model
        e, d1, d2 = m.enhance, m.denoise_1, m.denoise_2
        if True:
            fw = self.w
            self._wg = {
                "e.in": ops.WGrad(fw["e.in"], e.in_conv[0].weight.shape, device),
                "e.blk": ops.WGrad(fw["e.blk_raw"], e.conv[0].weight.shape, device),
                "e.out": ops.WGrad(fw["e.out"], e.out_conv[0].weight.shape, device),
            }
            for n, d in (("d1", d1), ("d2", d2)):
                for c in ("1", "2", "3"):
                    self._wg[f"{n}.c{c}"] = ops.WGrad(fw[f"{n}.c{c}"], getattr(d, "conv" + c).weight.shape, device)
            self._bn_g = torch.zeros(2, 64, device=device)

    def _denoise_bwd(self, n, tag, g_d, d_cs, H, W, cin_p, want_input_grad):
        """Backward of one application of denoise_1 / denoise_2 given the NHWC gradient of its output.
        Accumulates the weight gradients; returns the NHWC gradient of the packed input (denoise_2 only)."""
        ws, wb, wg, HW, T = self.ws, self._wbwd, self._wg, H * W, self.TC_BWD
        xin = ws.get(("xp" if n == "d1" else "dp12") + tag, (HW, cin_p))
        a1, a2 = ws.get("a48a" + tag, (HW, 48)), ws.get("a48b" + tag, (HW, 48))
        g2, g1 = ws.get("g48b", (HW, 48)), ws.get("g48a", (HW, 48))
        wg[n + ".c3"].accumulate(a2, 48, g_d, d_cs, H, W, (0, 0))
        ops.conv2d(wb[n + ".c3"], g_d, d_cs, d_cs, 1, H, W, g2, 48, tc=T)
        ops.act_bwd(g2, 48, a2, 48, 48, "lrelu", HW)
        wg[n + ".c2"].accumulate(a1, 48, g2, 48, H, W, (1, 1))
        ops.conv2d(wb[n + ".c2"], g2, 48, 48, 1, H, W, g1, 48, pad=(1, 1), tc=T)
        ops.act_bwd(g1, 48, a1, 48, 48, "lrelu", HW)
        wg[n + ".c1"].accumulate(xin, cin_p, g1, 48, H, W, (1, 1))
        if not want_input_grad:
            return None
        gin = ws.get("g12", (HW, cin_p))
        ops.conv2d(wb[n + ".c1"], g1, 48, 48, 1, H, W, gin, cin_p, pad=(1, 1), tc=T)
        return gin

    def backward(self):
        """Gradients of the summed loss w.r.t. the 20 trainable tensors (TRAIN_PARAMS order)."""
        c = self.ctx
        if c is None or "outs" not in c:
            raise RuntimeError("backward() needs a preceding forward_train() + loss_terms()")
        if not c["training"]:
            raise NotImplementedError("backward through eval-mode BatchNorm is not part of train.py's step")
        (L_pred1, L_pred2, L2, s2, s21, s22, H2, H11, H12, H13, s13, H14, s14, H3, s3, H3_pred, H4_pred,
         d12, dH3, H2_blur, H3_blur, H3d1, H3d2) = c["outs"]
        x, H, W = c["x"], c["H"], c["W"]
        h, wd = H // 2, W // 2
        HW, hw = H * W, h * wd
        n3, nh = 3.0 * HW, 3.0 * hw
        eps = self.EPS
        dev = x.device
        self._prepare_bwd(dev)
        ws, wg = self.ws, self._wg
        for g in wg.values():
            g.zero()
        self._bn_g.zero_()
        # frame-level gradient planes (zeroed slab; kernels accumulate)
        full = ws.get("gfull", (5, 1, 3, H, W))
        half = ws.get("ghalf", (12, 1, 3, h, wd))
        full.zero_()
        half.zero_()
        g_s2, g_H2, g_H3, g_s3, g_L2 = full[0], full[1], full[2], full[3], full[4]
        (g_Lp1, g_Lp2, g_H13, g_s13, g_H14, g_s14, g_H3d1, g_H3d2, g_H11, g_H12, g_s21, g_s22) = half.unbind(0)
        tf = [ws.get("bw_tf%d" % i, (1, 3, H, W)) for i in range(3)]
        th = [ws.get("bw_th%d" % i, (1, 3, h, wd)) for i in range(2)]
        # ---- loss adjoints (loss.py:46-77) ----
        ops.loss_illum_bwd(L2, s2, c["is_WB"], ws.get("chansum", (4,), torch.float64), 700 * 2 / n3, 1000 * 2 / n3, g_s2)
        ops.loss_smooth_bwd(ws.get("loss_ycc", (1, 3, H, W)), s2,
                            [5 * 2.0 / ((H - dy) * (W - abs(dx))) for dy, dx in SMOOTH_OFFSETS], g_s2)
        ops.loss_tv_bwd(s2, 1600 * 2.0 / ((H - 1) * W), 1600 * 2.0 / (H * (W - 1)), g_s2)
        k1 = 1000 * 2 / nh
        ops.mse_bwd(c["L11e"], L_pred2, k1, gb=g_Lp2)
        ops.mse_bwd(c["L12e"], L_pred1, k1, gb=g_Lp1)
        th[0].zero_()
        th[1].zero_()
        ops.mse_bwd(L_pred1, c["den1"], k1, ga=g_Lp1, gb=th[0])
        ops.mse_bwd(L_pred2, c["den2"], k1, ga=g_Lp2, gb=th[1])
        ops.pair_downsample_bwd(th[0], th[1], g_L2)
        k2 = 1000 * 2 / (2 * nh)
        ops.mse_bwd(H13, H12, k2, ga=g_H13)
        ops.mse_bwd(s13, s22, k2, ga=g_s13)
        ops.mse_bwd(H14, H11, k2, ga=g_H14)
        ops.mse_bwd(s14, s21, k2, ga=g_s14)
        ops.mse_bwd(H13, H3d1, k1, ga=g_H13, gb=g_H3d1)
        ops.mse_bwd(H14, H3d2, k1, ga=g_H14, gb=g_H3d2)
        tf[0].zero_()
        ops.mse_bwd(H3_blur, H2_blur, 10000 * 2 / n3, ga=tf[0])
        ops.blur21_bwd(tf[0], tf[1], g_H3)
        ops.mse_bwd(s3, s2, 1000 * 2 / n3, ga=g_s3)
        ops.loss_inter_bwd(H3d1, H3d2, dH3, 10000 * 2 / nh, th[0], th[1], g_H3d1, g_H3d2)
        ops.loss_var_bwd(H2, H3, ws.get("loss_avg1", (1, 3, H, W)), ws.get("loss_avg2", (1, 3, H, W)), 1000 * 2 / n3,
                         tf[0], tf[1], tf[2], g_H2, g_H3)
        ops.pair_downsample_bwd(g_H3d1, g_H3d2, g_H3)
        # ---- denoise_2, three applications (model.py:180-191); inputs H*, s* are NOT detached ----
        for tag, hh, ww, gH, oH, gs, os_, dH, ds in ((".f", H, W, g_H3, H3, g_s3, s3, g_H2, g_s2),
                                                     (".a", h, wd, g_H13, H13, g_s13, s13, g_H11, g_s21),
                                                     (".b", h, wd, g_H14, H14, g_s14, s14, g_H12, g_s22)):
            gd = ws.get("gd8", (hh * ww, 8))
            ops.mask_pack([gH, gs], [oH, os_], 3, -1.0, eps, 1.0, gd, 8, hh * ww)
            gin = self._denoise_bwd("d2", tag, gd, 8, hh, ww, 12, True)
            ops.unpack_add(gin, 12, 6, 3, [dH, ds], hh * ww)
        # ---- Retinex glue (model.py:169-177) ----
        ops.div_clamp_bwd(c["L11"], 0.0, s21, g_H11, eps, 1.0, g_s21)
        ops.div_clamp_bwd(c["L12"], 0.0, s22, g_H12, eps, 1.0, g_s22)
        ops.pair_downsample_bwd(g_s21, g_s22, g_s2)
        ops.div_clamp_bwd(x, eps, s2, g_H2, eps, 1.0, g_s2)
        # ---- enhancer (model.py:74-81), BatchNorm in train mode, one shared block applied three times ----
        bn = self.model.enhance.conv[1]
        T = self.TC_BWD
        s4 = ws.get("s4", (HW, 4))
        cur = [ws.get("e.cur%d" % k, (HW, 64)) for k in range(4)]
        t = [ws.get("e.t%d" % k, (HW, 64)) for k in range(3)]
        st = [ws.get("e.stats%d" % k, (128,), torch.float64) for k in range(3)]
        red = ws.get("bn_red", (128,), torch.float64)
        gs4 = ws.get("gs4", (HW, 4))
        ga, gb, gt = ws.get("g64a", (HW, 64)), ws.get("g64b", (HW, 64)), ws.get("g64t", (HW, 64))
        ops.mask_pack([g_s2], [s2], 3, 1.0, 1e-4, 1.0, gs4, 4, HW)   # Enhancer clamps to [1e-4, 1] (model.py:80)
        ops.act_bwd(gs4, 4, s4, 4, 4, "sigmoid", HW)
        wg["e.out"].accumulate(cur[3], 64, gs4, 4, H, W, (1, 1))
        ops.conv2d(self._wbwd["e.out"], gs4, 4, 4, 1, H, W, ga, 64, pad=(1, 1), tc=T)
        dcur, dnxt = ga, gb
        for k in (2, 1, 0):
            ops.bn_train_bwd(t[k], 64, dcur, 64, st[k], red, bn.weight, bn.bias, HW, 64, bn.eps, gt, 64,
                             self._bn_g[0], self._bn_g[1])
            wg["e.blk"].accumulate(cur[k], 64, gt, 64, H, W, (1, 1))
            ops.conv2d(self._wbwd["e.blk"], gt, 64, 64, 1, H, W, dnxt, 64, pad=(1, 1), epi="add", aux0=dcur,
                       aux0_cs=64, tc=T)
            dcur, dnxt = dnxt, dcur
        ops.act_bwd(dcur, 64, cur[0], 64, 64, "relu", HW)
        wg["e.in"].accumulate(ws.get("ep12", (HW, 12)), 12, dcur, 64, H, W, (1, 1))
        # ---- denoise_1, three applications (model.py:149-152) ----
        gd = ws.get("gd4", (HW, 4))
        ops.mask_pack([g_L2], [L2], 3, -1.0, eps, 1.0, gd, 4, HW)
        self._denoise_bwd("d1", ".1f", gd, 4, H, W, 4, False)
        for tag, g in ((".1a", g_Lp1), (".1b", g_Lp2)):
            gd = ws.get("gd4h", (hw, 4))
            ops.mask_pack([g], [None], 3, -1.0, 0.0, 0.0, gd, 4, hw)
            self._denoise_bwd("d1", tag, gd, 4, h, wd, 4, False)
        # ---- parameter gradients in torch layout ----
        out = []
        for name in ("e.in", "e.blk"):
            out += list(wg[name].grads())
            if name == "e.blk":
                out += [self._bn_g[0].clone(), self._bn_g[1].clone()]
        out += list(wg["e.out"].grads())
        for n in ("d1", "d2"):
            for cc in ("1", "2", "3"):
                out += list(wg[f"{n}.c{cc}"].grads())
        return out

"""Thin torch-tensor wrappers over the libzerotig C ABI (include/zerotig.h).

Each wrapper borrows ``data_ptr()``s for the duration of one call on torch's current
CUDA stream; outputs are caller- or torch-allocated.  No wrapper has a CPU path: a
non-CUDA tensor raises.
"""
import ctypes as C
import math

import torch

from . import _lib
from ._lib import ZtConv, ZtConvBf16, ZtRaftUpdate, check, lib

F16_ACT_SCALE = 16.0      # ZT_F16_ACT_SCALE of include/zerotig.h (split mode 4)
ACT = {"none": 0, "relu": 1, "lrelu": 2, "sigmoid": 3, "tanh": 4}
EPI = {"plain": 0, "gru_zr": 1, "gru_q": 2, "add_relu": 3, "add": 4}


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t, offset_elems=0):
    """Device pointer (as c_void_p) of a float tensor, optionally offset by elements."""
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise RuntimeError("zero_tig_b200 ops need CUDA tensors (no CPU fallback)")
    return C.c_void_p(t.data_ptr() + offset_elems * t.element_size())


class ConvW:
    """A conv layer's parameters in kernel layout: weight [kh][kw][cin_pad][w_cs], bias [cout]."""
    __slots__ = ("w", "b", "kh", "kw", "cin", "cout", "w_cs", "w_raw", "b_raw", "_tc", "w_scale")

    def __init__(self, weight, bias=None, cin_pad=None, bn=None, eps=1e-5, out_slice=None):
        w = weight.detach().float()
        b = None if bias is None else bias.detach().float()
        if bn is not None:  # fold eval-mode BatchNorm: (gamma, beta, mean, var)
            g, beta, mean, var = [t.detach().float() for t in bn]
            s = g / torch.sqrt(var + eps)
            w = w * s.view(-1, 1, 1, 1)
            b0 = b if b is not None else torch.zeros_like(mean)
            b = (b0 - mean) * s + beta
        if out_slice is not None:
            w = w[out_slice]
            b = None if b is None else b[out_slice]
        cout, cin, kh, kw = w.shape
        cin_p = cin_pad or ((cin + 3) // 4 * 4)
        w_cs = (cout + 3) // 4 * 4
        wk = torch.zeros(kh, kw, cin_p, w_cs, device=w.device, dtype=torch.float32)
        wk[:, :, :cin, :cout] = w.permute(2, 3, 1, 0)
        self.w, self.b = wk.contiguous(), (None if b is None else b.contiguous())
        self.kh, self.kw, self.cin, self.cout, self.w_cs = kh, kw, cin_p, cout, w_cs
        self.w_raw, self._tc = w, {}
        self.b_raw = b
        self.w_scale = None      # power of two applied to the weights of the scaled-fp16 packing (split mode 4)

    def f16_scale(self):
        """Power of two that brings max|w| into (4096, 8192] (split mode 4): the fp16 low halves of the scaled weights
        stay normal numbers (>= 2^-14), so hi + lo carries ~22 bits.  Computed once per ConvW (one host sync)."""
        if self.w_scale is None:
            m = float(self.w_raw.abs().max())
            e = 13 - math.frexp(m)[1] if m > 0 else 10        # m = f * 2^e', f in [0.5, 1): m * 2^(13 - e') in [4096, 8192)
            self.w_scale = float(2.0 ** max(min(e, 24), -8))
        return self.w_scale

    def acc_scale(self):
        """ZtConv.acc_scale of split mode 4: undoes the activation and weight scales on the fp32 accumulator."""
        return 1.0 / (F16_ACT_SCALE * self.f16_scale())

    def tc_pack(self, BN, fp16=False):
        """hi/lo weights in the tcgen05 kernel's layout (csrc/conv_tc.cu):
        [n_tile][64-channel group][tap][hi|lo][row BN][128 B], each 128 B row (64 input channels
        of one tap) stored with the UMMA 128-byte swizzle (16 B chunk c of row n at chunk
        position c ^ (n & 7)) - one contiguous block per (n_tile, group, tap) = one
        cp.async.bulk per weight stage.  bf16 pairs of w (split modes 1, 3) or fp16 pairs of w * f16_scale() (mode 4)."""
        key = (BN, "f16") if fp16 else BN
        t = self._tc.get(key)
        if t is None:
            w = self.w_raw
            cout, cin, kh, kw = w.shape
            ng, nt = (self.cin + 63) // 64, (cout + BN - 1) // BN
            wp = torch.zeros(nt * BN, ng * 64, kh, kw, device=w.device, dtype=torch.float32)
            wp[:cout, :cin] = w
            dt = torch.float16 if fp16 else torch.bfloat16
            if fp16:
                wp = wp * self.f16_scale()
            hi = wp.to(dt)
            lo = (wp - hi.float()).to(dt)
            x = torch.stack([hi, lo], 0).view(torch.int16).reshape(2, nt, BN, ng, 8, 8, kh * kw)
            x = x.permute(1, 3, 6, 0, 2, 4, 5).contiguous()      # nt, g, tap, part, row, chunk, e
            n = torch.arange(BN, device=w.device)
            src = torch.arange(8, device=w.device)[None, :] ^ (n[:, None] & 7)   # position j <- chunk j^(n&7)
            idx = src.view(1, 1, 1, 1, BN, 8, 1).expand_as(x)
            t = torch.gather(x, 5, idx).contiguous()
            # constant weights are fetched BEFORE griddepcontrol.wait by the kernel (programmatic dependent
            # launch): make sure the one-time packing has really finished before the first launch
            torch.cuda.current_stream().synchronize()
            self._tc[key] = t
        return t


    def tma_pack(self):
        """bf16 weights for csrc/conv_tma.cu: [tap][BN = ceil16(cout)][64 ch] rows of 128 B, 16 B chunk j of
        row n stored at chunk position j ^ (n & 7) (UMMA / TMA 128-byte swizzle); cin, cout zero padded."""
        t = self._tc.get("tma")
        if t is None:
            w = self.w_raw
            cout, cin, kh, kw = w.shape
            assert cin <= 64 and cout <= 64
            BN = (cout + 15) // 16 * 16
            wp = torch.zeros(BN, 64, kh, kw, device=w.device, dtype=torch.float32)
            wp[:cout, :cin] = w
            x = wp.to(torch.bfloat16).permute(2, 3, 0, 1).reshape(kh * kw, BN, 8, 8)   # tap, row, chunk, e
            n = torch.arange(BN, device=w.device)
            src = torch.arange(8, device=w.device)[None, :] ^ (n[:, None] & 7)
            t = torch.gather(x, 2, src.view(1, BN, 8, 1).expand_as(x)).contiguous()
            torch.cuda.current_stream().synchronize()   # fetched before griddepcontrol.wait (see tc_pack)
            self._tc["tma"] = t
        return t


    def head_pack(self):
        """fp32 [8][64] weights + [8] bias of a 1x1 convolution fused into the preceding layer's epilogue
        (csrc/conv_tma.cu, OUT = 3); rows / columns beyond cout / cin are zero."""
        t = self._tc.get("head")
        if t is None:
            w = self.w_raw
            cout, cin, kh, kw = w.shape
            assert kh == 1 and kw == 1 and cout <= 8 and cin <= 64
            hw = torch.zeros(8, 64, device=w.device, dtype=torch.float32)
            hw[:cout, :cin] = w[:, :, 0, 0]
            hb = torch.zeros(8, device=w.device, dtype=torch.float32)
            if self.b_raw is not None:
                hb[:cout] = self.b_raw
            torch.cuda.current_stream().synchronize()   # fetched before griddepcontrol.wait (see tc_pack)
            t = (hw, hb)
            self._tc["head"] = t
        return t


def conv2d_bf16(wt: ConvW, x, cin, H, W, out, out_cs, *, act="none", res=None, res_cs=0, out_f32=False, head=None):
    """TMA-fed tcgen05 convolution over bf16 NHWC activations (csrc/conv_tma.cu): 3x3 pad 1 or 1x1.
    head = ConvW of a following 1x1 convolution (cout <= 8) applied in the epilogue: out is then its fp32 result."""
    d = ZtConvBf16()
    if head is not None:
        hw, hb = head.head_pack()
        d.head_w, d.head_b, d.head_n = _p(hw), _p(hb), head.w_raw.shape[0]
        out_f32 = True
    d.in_, d.cin = _p(x), cin
    d.weight, d.bias = _p(wt.tma_pack()), _p(wt.b_raw)
    d.out, d.out_cs, d.out_f32 = _p(out), out_cs, int(out_f32)
    d.res, d.res_cs = _p(res), res_cs
    d.H, d.W, d.cout, d.kh, d.kw, d.act = H, W, wt.w_raw.shape[0], wt.kh, wt.kw, ACT[act]
    check(lib().zt_conv2d_bf16_tma(C.byref(d), _stream()), "conv2d_bf16_tma")


def pack_nhwc_bf16(srcs, c_each, add, out, cs, HW):
    s = list(srcs) + [None] * (4 - len(srcs))
    check(lib().zt_pack_nhwc_bf16(_p(s[0]), _p(s[1]), _p(s[2]), _p(s[3]), c_each, len(srcs), add, _p(out), cs, HW,
                                  _stream()), "pack_nhwc_bf16")


def conv2d(wt: ConvW, in0, in0_cs, c0, B, H, W, out, out_cs, *, in0_off=0, in1=None, in1_cs=0, c1=0,
           in1_off=0, out_off=0, stride=1, pad=(0, 0), act="none", scale=1.0, epi="plain",
           aux0=None, aux0_cs=0, aux0_off=0, aux1=None, aux1_cs=0, aux1_off=0, tc=0, bn=None, crop=None, bplane=None, bplane_cs=0):
    """out[:, out_off:out_off+cout] = epilogue(conv(cat[in0[:, in0_off:+c0], in1[:, in1_off:+c1]])).
    crop = (Ho, Wo) keeps only the top-left part of the output (tensor-core path; see s2d_weight)."""
    assert c0 + c1 == wt.cin, (c0, c1, wt.cin)
    sh = sw = stride
    ph, pw = pad
    Ho = (H + 2 * ph - wt.kh) // sh + 1
    Wo = (W + 2 * pw - wt.kw) // sw + 1
    if crop is not None:
        assert tc and crop[0] <= Ho and crop[1] <= Wo
        Ho, Wo = crop
    d = ZtConv()
    d.in0, d.in0_cs, d.c0 = _p(in0, in0_off), in0_cs, c0
    d.in1, d.in1_cs, d.c1 = _p(in1, in1_off), in1_cs, c1
    d.weight, d.w_cs = _p(wt.w), wt.w_cs
    d.bias = _p(wt.b)
    d.out, d.out_cs = _p(out, out_off), out_cs
    d.B, d.H, d.W, d.Ho, d.Wo, d.cout = B, H, W, Ho, Wo, wt.cout
    d.kh, d.kw, d.sh, d.sw, d.ph, d.pw = wt.kh, wt.kw, sh, sw, ph, pw
    d.act, d.out_scale, d.epi = ACT[act], scale, EPI[epi]
    d.aux0, d.aux0_cs = _p(aux0, aux0_off), aux0_cs
    d.aux1, d.aux1_cs = _p(aux1, aux1_off), aux1_cs
    d.bplane, d.bplane_cs = _p(bplane), bplane_cs
    if tc:
        # tensor-core path: tc = 1 (bf16), 3 (bf16 hi/lo split, 3 products) or 4 (scaled fp16 hi/lo split, 3 products:
        # fp32-grade)
        BN = bn or min((wt.cout + 15) // 16 * 16, 128)
        if tc == 4:
            d.weight = _p(wt.tc_pack(BN, fp16=True))
            d.acc_scale = wt.acc_scale()
        else:
            d.weight = _p(wt.tc_pack(BN))
        # bit 8: weights written by the preceding kernel (RowsAsWeights) - see the PDL note in csrc/conv_tc.cu
        check(lib().zt_conv2d_tc_nhwc(C.byref(d), BN, tc | (0x100 if getattr(wt, "dynamic", False) else 0), _stream()),
              "conv2d_tc")
    else:
        check(lib().zt_conv2d_nhwc_f32(C.byref(d), _stream()), "conv2d")
    return Ho, Wo


class DevConvW:
    """Conv weights that are (re)packed ON THE DEVICE from the live nn.Parameter (zt_pack_tc_weights_f32): the
    zero-shot training loop (train.py:126-131) changes the weights every step, and packing them with host tensor ops
    cost ~20 small launches plus a stream synchronise per layer and step.  ``repack()`` after every optimiser step; the
    convolution is launched with the dynamic-weights flag so it waits for the packer.  dgrad=True packs the weights of
    the data-gradient convolution (channels swapped, taps mirrored).  Tensor-core modes only (3 = bf16 pairs, 4 =
    scaled fp16 pairs with a fixed weight scale: trainable weights start at N(0, 0.02) and stay far below 64)."""
    dynamic = True
    F16_W_SCALE = 1024.0

    def __init__(self, weight, bias=None, cin_pad=None, dgrad=False):
        self.weight, self.b, self.dgrad = weight, bias, dgrad
        co, ci, self.kh, self.kw = weight.shape
        self.cout, cin = (ci, co) if dgrad else (co, ci)
        self.cin = cin_pad or ((cin + 3) // 4 * 4)
        self.w_cs = (self.cout + 3) // 4 * 4
        self.w, self.w_scale, self._bufs = None, None, {}
        self.w_raw = weight.detach().transpose(0, 1) if dgrad else weight.detach()     # shape bookkeeping (bench, WGrad)
        self.b_raw = bias

    def f16_scale(self):
        return self.F16_W_SCALE

    def acc_scale(self):
        return 1.0 / (F16_ACT_SCALE * self.F16_W_SCALE)

    def _pack(self, BN, fp16, buf):
        w = self.weight.detach()
        co, ci = w.shape[0], w.shape[1]
        check(lib().zt_pack_tc_weights_f32(_p(w), co, ci, self.kh, self.kw, self.cin, int(self.dgrad), BN, int(fp16),
                                           self.F16_W_SCALE if fp16 else 1.0, _p(buf), _stream()), "pack_tc_weights")

    def tc_pack(self, BN, fp16=False):
        key = (BN, bool(fp16))
        buf = self._bufs.get(key)
        if buf is None:
            nt, ng = (self.cout + BN - 1) // BN, (self.cin + 63) // 64
            buf = torch.empty(nt * ng * self.kh * self.kw * 2 * BN * 128, dtype=torch.uint8, device=self.weight.device)
            self._bufs[key] = buf
            self._pack(BN, fp16, buf)
        return buf

    def repack(self):
        for (BN, fp16), buf in self._bufs.items():
            self._pack(BN, fp16, buf)


class RowsAsWeights:
    """Device rows [R][C] used as the weights of a 1x1 tensor-core convolution with cout = R (the
    all-pairs correlation corr.py:52-60: weights = fmap2, re-packed on the device every frame)."""
    dynamic = True

    def __init__(self, R, Cn, BN, device):
        self.kh = self.kw = 1
        self.cin, self.cout, self.b, self.BN = Cn, R, None, BN
        self.w, self.w_cs = None, 0
        nt, ng = (R + BN - 1) // BN, (Cn + 63) // 64
        self.buf = torch.empty(nt * ng * 2 * BN * 128, dtype=torch.uint8, device=device)

    def pack(self, rows, rs):
        check(lib().zt_pack_tc_rows_f32(_p(rows), self.cout, self.cin, rs, self.BN, _p(self.buf), _stream()),
              "pack_tc_rows")

    def tc_pack(self, BN):
        assert BN == self.BN
        return self.buf


def s2d_weight(w, pad):
    """Weights of the stride-1 convolution over the space-to-depth input that equals a stride-2 convolution
    (square kernel k, padding `pad`): w [co][ci][k][k] -> ([co][4*ci][n][n], new_pad) with input row
    2*(oy+a)+b <-> tap a, block row b (same for columns); taps without a source are zero."""
    co, ci, k, _ = w.shape
    amin, amax = (0 - pad) // 2, (k - 1 - pad) // 2
    n = amax - amin + 1
    out = torch.zeros(co, 4 * ci, n, n, device=w.device, dtype=w.dtype)
    for ky in range(k):
        ay, by = divmod(ky - pad, 2)
        for kx in range(k):
            ax, bx = divmod(kx - pad, 2)
            blk = by * 2 + bx
            out[:, blk * ci:(blk + 1) * ci, ay - amin, ax - amin] = w[:, :, ky, kx]
    return out, -amin


def space_to_depth2(x, x_cs, Cn, B, H, W, out):
    check(lib().zt_space_to_depth2_f32(_p(x), x_cs, Cn, B, H, W, _p(out), _stream()), "space_to_depth2")


def raft_convf1(flow, f_cs, w, b, out, out_cs, H, W):
    check(lib().zt_raft_convf1_f32(_p(flow), f_cs, _p(w), _p(b), _p(out), out_cs, H, W, _stream()), "raft_convf1")


def conv3x3_n2(x, x_cs, Cn, w, b, out, out_cs, H, W):
    check(lib().zt_conv3x3_n2_f32(_p(x), x_cs, Cn, _p(w), _p(b), _p(out), out_cs, H, W, _stream()), "conv3x3_n2")


def flow_head2_update(x, x_cs, Cn, w, b, coords1, f1, f1_cs, f2, f2_cs, f2_off, H, W):
    """FlowHead.conv2 + coords1 += delta + flow outputs in one launch (update.py:13-14, raft.py:117-118)."""
    check(lib().zt_flow_head2_update_f32(_p(x), x_cs, Cn, _p(w), _p(b), _p(coords1), _p(f1), f1_cs, _p(f2, f2_off), f2_cs,
                                         H, W, _stream()), "flow_head2_update")


RAFT_UPDATE_LAYERS = ("convc1", "convf2", "convc2", "conv", "zr1", "q1", "zr2", "q2", "fh1")


def raft_update(h8, w8, iters, bufs, bplanes, levels, hs, ws, weights, bns, wf1, bf1, wfh2, bfh2, max_ctas=0, fp16=False):
    """All `iters` iterations of the RAFT update block in one persistent cooperative kernel (csrc/raft_update.cu;
    raft.py:111-120).  bufs: dict of the device buffers named in ZtRaftUpdate (corr_hi .. rh_lo bf16, h / z / fhid /
    coords1 / flow4 fp32, counter int32); bplanes: (zr1, q1, zr2, q2); weights: ConvW per RAFT_UPDATE_LAYERS entry,
    bns: their N tiles."""
    d = ZtRaftUpdate()
    d.h8, d.w8, d.iters, d.max_ctas = h8, w8, iters, max_ctas
    for n in ("corr", "c1", "f1", "c2f", "hm", "rh"):
        for sfx in ("_hi", "_lo"):
            setattr(d, n + sfx, _p(bufs[n + sfx]))
    for n in ("h", "z", "fhid", "coords1", "flow4", "counter"):
        setattr(d, n, _p(bufs[n]))
    d.bp_zr1, d.bp_q1, d.bp_zr2, d.bp_q2 = [_p(t) for t in bplanes]
    for i in range(4):
        d.lvl[i], d.lh[i], d.lw[i] = levels[i].data_ptr(), hs[i], ws[i]
    d.lvl_bf16 = int(levels[0].dtype == torch.bfloat16)
    d.fp16 = int(fp16)
    for i, (wt, bn) in enumerate(zip(weights, bns)):
        d.w[i] = wt.tc_pack(bn, fp16=fp16).data_ptr()
        d.b[i] = wt.b.data_ptr() if wt.b is not None else None
        d.bn[i] = bn
        d.acc_scale[i] = wt.acc_scale() if fp16 else 1.0
    d.wf1, d.bf1, d.wfh2, d.bfh2 = _p(wf1), _p(bf1), _p(wfh2), _p(bfh2)
    check(lib().zt_raft_update(C.byref(d), _stream()), "raft_update")


def split_bf16_hilo(x, rs, hi, lo, out_rs, Cn, N):
    check(lib().zt_split_bf16_hilo(_p(x), rs, _p(hi), _p(lo), out_rs, Cn, N, _stream()), "split_bf16_hilo")


def instnorm(x, x_cs, out, out_cs, stats, B, HW, Cn, relu_inner, relu_outer=False, res=None, res_cs=0, eps=1e-5):
    check(lib().zt_instnorm_nhwc_f32(_p(x), x_cs, _p(out), out_cs, _p(res), res_cs, _p(stats), B, HW, Cn,
                                     eps, int(relu_inner), int(relu_outer), _stream()), "instnorm")


def transpose(x, R, Cc, in_rs, out, out_rs):
    check(lib().zt_transpose_f32(_p(x), R, Cc, in_rs, _p(out), out_rs, _stream()), "transpose")


def avgpool2(x, out, planes, h, w):
    check(lib().zt_avgpool2_f32(_p(x), _p(out), planes, h, w, _stream()), "avgpool2")


def corr_pyramid_scratch_bytes(h0, w0):
    return int(lib().zt_corr_pyramid_scratch_bytes(h0, w0))


def corr_pyramid(fmap, fm_rs, h0, w0, levels, scratch):
    """All-pairs correlation + 4-level pyramid in one tcgen05 kernel (corr.py:13-27,52-60).  The storage type of
    `levels` (fp32 or bf16) selects the entry point."""
    fn = lib().zt_corr_pyramid_bf16 if levels[0].dtype == torch.bfloat16 else lib().zt_corr_pyramid_f32
    check(fn(_p(fmap), fm_rs, h0, w0, _p(levels[0]), _p(levels[1]), _p(levels[2]),
                                    _p(levels[3]), _p(scratch), scratch.numel() * scratch.element_size(), _stream()),
          "corr_pyramid")


def corr_lookup(levels, hs, ws, coords, out, out_cs, N):
    IA = C.c_int * 4
    fn = lib().zt_corr_lookup_bf16 if levels[0].dtype == torch.bfloat16 else lib().zt_corr_lookup_f32
    check(fn(_p(levels[0]), _p(levels[1]), _p(levels[2]), _p(levels[3]),
             IA(*hs), IA(*ws), _p(coords), _p(out), out_cs, N, _stream()), "corr_lookup")


def convex_upsample(flow, flow_cs, mask, mask_cs, out, h8, w8):
    check(lib().zt_convex_upsample_f32(_p(flow), flow_cs, _p(mask), mask_cs, _p(out), h8, w8, _stream()),
          "convex_upsample")


def flow_update(coords1, delta, d_cs, f1, f1_cs, f2, f2_cs, f2_off, h8, w8):
    check(lib().zt_flow_update_f32(_p(coords1), _p(delta), d_cs, _p(f1), f1_cs, _p(f2, f2_off), f2_cs, h8, w8,
                                   _stream()), "flow_update")


def resize255_image1(img, H, W, ht, wd, out, hp, wp):
    check(lib().zt_resize255_f32(_p(img), H, W, ht, wd, 0, _p(out), hp, wp, None, None, _stream()), "resize255")


def resize255_quant_hist(img, H, W, ht, wd, out_u8, hist):
    check(lib().zt_resize255_f32(_p(img), H, W, ht, wd, 1, None, ht, wd, _p(out_u8), _p(hist), _stream()),
          "resize255")


def equalize_lut(hist, lut):
    check(lib().zt_equalize_lut(_p(hist), _p(lut), _stream()), "equalize_lut")


def equalize_apply_pack(q, lut, ht, wd, hp, wp, out, eq_u8=None):
    check(lib().zt_equalize_apply_pack(_p(q), _p(lut), ht, wd, hp, wp, _p(out), _p(eq_u8), _stream()),
          "equalize_apply_pack")


def warp_pair(flow, hf, wf, img_a, img_b, out_a, out_b, Cn, H, W):
    check(lib().zt_warp_pair_f32(_p(flow), hf, wf, _p(img_a), _p(img_b), _p(out_a), _p(out_b), Cn, H, W,
                                 _stream()), "warp_pair")


def pack_nhwc(srcs, c_each, add, out, cs, HW):
    s = list(srcs) + [None] * (4 - len(srcs))
    check(lib().zt_pack_nhwc_f32(_p(s[0]), _p(s[1]), _p(s[2]), _p(s[3]), c_each, len(srcs), add, _p(out), cs, HW,
                                 _stream()), "pack_nhwc")


def retinex_l2(x, d, d_cs, L2, HW, eps):
    check(lib().zt_retinex_l2_f32(_p(x), _p(d), d_cs, _p(L2), HW, eps, _stream()), "retinex_l2")


def retinex_h2(x, s, s_cs, s2, H2, HW, eps):
    check(lib().zt_retinex_h2_f32(_p(x), _p(s), s_cs, _p(s2), _p(H2), HW, eps, _stream()), "retinex_h2")


def retinex_h5(H2, s2, d, d_cs, H3, s3, HW, eps):
    check(lib().zt_retinex_h5_f32(_p(H2), _p(s2), _p(d), d_cs, _p(H3), _p(s3), HW, eps, _stream()), "retinex_h5")


def quantize_u8_hwc(x, out, Cn, HW):
    check(lib().zt_quantize_u8_hwc(_p(x), _p(out), Cn, HW, _stream()), "quantize_u8")


def u8_hwc_to_f32(img_u8, out, Cn, HW):
    """transforms.ToTensor on the device: uint8 HWC -> float NCHW / 255 (multi_read_data.py:86-87,131)."""
    check(lib().zt_u8_hwc_to_f32(_p(img_u8), _p(out), Cn, HW, _stream()), "u8_hwc_to_f32")


# ---------------------------------------------------------------------------------------
# training path (csrc/train.cu, csrc/norm_f32.cu)
# ---------------------------------------------------------------------------------------
def bn_train(x, x_cs, out, out_cs, res, res_cs, stats, gamma, beta, rmean, rvar, HW, Cn, eps, momentum):
    check(lib().zt_bn_train_nhwc_f32(_p(x), x_cs, _p(out), out_cs, _p(res), res_cs, _p(stats), _p(gamma), _p(beta),
                                     _p(rmean), _p(rvar), HW, Cn, eps, momentum, _stream()), "bn_train")


def pair_downsample(x, add=0.0):
    """utils.pair_downsampler on an NCHW [1,C,H,W] tensor (+add) -> two [1,C,H/2,W/2]."""
    _, Cn, H, W = x.shape
    o1 = torch.empty(1, Cn, H // 2, W // 2, device=x.device)
    o2 = torch.empty_like(o1)
    check(lib().zt_pair_downsample_f32(_p(x), _p(o1), _p(o2), Cn, H, W, add, _stream()), "pair_downsample")
    return o1, o2


def sub_clamp(a, add, d, d_cs, out, Cn, HW, lo=0.0, hi=0.0, clamp=False):
    check(lib().zt_sub_clamp_f32(_p(a), add, _p(d), d_cs, _p(out), Cn, HW, lo, hi, int(clamp), _stream()), "sub_clamp")


def div_clamp(a, add, b, lo, hi):
    out = torch.empty_like(a)
    check(lib().zt_div_clamp_f32(_p(a), add, _p(b), _p(out), a.numel(), lo, hi, _stream()), "div_clamp")
    return out


def texture_diff(i1, i2):
    _, _, H, W = i1.shape
    out = torch.empty(1, 1, H, W, device=i1.device)
    check(lib().zt_texture_diff_f32(_p(i1), _p(i2), _p(out), H, W, _stream()), "texture_diff")
    return out


_BLUR_TAPS = None


def blur_taps():
    """1-D taps of utils.blur (utils.py:29-39 with nsig=1): sqrt(k)/sum(sqrt(k)), k = diff of the normal
    CDF on 22 points - the reference's 2-D kernel sqrt(outer(k,k))/sum is exactly their outer product."""
    global _BLUR_TAPS
    if _BLUR_TAPS is None:
        import math
        n, nsig = 21, 1
        interval = (2 * nsig + 1.0) / n
        lo, hi = -nsig - interval / 2.0, nsig + interval / 2.0
        xs = [lo + (hi - lo) * i / n for i in range(n + 1)]
        cdf = [0.5 * (1 + math.erf(v / math.sqrt(2.0))) for v in xs]
        r = [math.sqrt(cdf[i + 1] - cdf[i]) for i in range(n)]
        t = sum(r)
        _BLUR_TAPS = (C.c_float * 21)(*[v / t for v in r])
    return _BLUR_TAPS


def blur21(x):
    _, Cn, H, W = x.shape
    tmp, out = torch.empty_like(x), torch.empty_like(x)
    check(lib().zt_blur21_f32(_p(x), _p(tmp), _p(out), blur_taps(), Cn, H, W, _stream()), "blur21")
    return out


def loss_illum(L2, s2, is_WB, chansum, acc, off):
    check(lib().zt_loss_illum_f32(_p(L2), _p(s2), int(is_WB), L2.shape[2] * L2.shape[3], _p(chansum), _p(acc, off),
                                  _stream()), "loss_illum")


def loss_smooth(L2, s2, ycc, acc, off):
    check(lib().zt_loss_smooth_f32(_p(L2), _p(s2), _p(ycc), L2.shape[2], L2.shape[3], _p(acc, off), _stream()),
          "loss_smooth")


def loss_tv(x, acc, off):
    check(lib().zt_loss_tv_f32(_p(x), x.shape[1], x.shape[2], x.shape[3], _p(acc, off), _stream()), "loss_tv")


def mse_sum(a, b, acc, off, add_a=0.0, n=None):
    check(lib().zt_mse_sum_f32(_p(a), _p(b), add_a, n if n is not None else a.numel(), _p(acc, off), _stream()), "mse_sum")


def loss_inter(d1, d2, diff, acc, off):
    check(lib().zt_loss_inter_f32(_p(d1), _p(d2), _p(diff), d1.shape[1], d1.shape[2], d1.shape[3], _p(acc, off),
                                  _stream()), "loss_inter")


def loss_var(H2, H3, avg1, avg2, acc, off):
    check(lib().zt_loss_var_f32(_p(H2), _p(H3), _p(avg1), _p(avg2), H2.shape[1], H2.shape[2], H2.shape[3],
                                _p(acc, off), _stream()), "loss_var")


# ---------------------------------------------------------------------------------------
# training backward (csrc/train_bwd.cu)
# ---------------------------------------------------------------------------------------
def dgrad_weights(weight, cin_pad=None):
    """Weights of the data-gradient convolution of a stride-1 nn.Conv2d: spatially flipped, in/out
    channels swapped, so dX = conv2d(dY, W') runs on the forward (tensor-core) convolution."""
    return ConvW(weight.detach().transpose(0, 1).flip(2, 3).contiguous(), None, cin_pad=cin_pad)


def mse_bwd(a, b, coef, ga=None, gb=None, add_a=0.0):
    check(lib().zt_mse_bwd_f32(_p(a), _p(b), add_a, a.numel(), coef, _p(ga), _p(gb), _stream()), "mse_bwd")


def loss_illum_bwd(L2, s2, is_WB, chansum, c0, c1, g_s2):
    check(lib().zt_loss_illum_bwd_f32(_p(L2), _p(s2), int(is_WB), L2.shape[2] * L2.shape[3], _p(chansum), c0, c1,
                                      _p(g_s2), _stream()), "loss_illum_bwd")


def loss_smooth_bwd(ycc, s2, coefs, g_s2):
    check(lib().zt_loss_smooth_bwd_f32(_p(ycc), _p(s2), s2.shape[2], s2.shape[3], (C.c_float * 12)(*coefs), _p(g_s2),
                                       _stream()), "loss_smooth_bwd")


def loss_tv_bwd(x, ch, cw, g):
    check(lib().zt_loss_tv_bwd_f32(_p(x), x.shape[1], x.shape[2], x.shape[3], ch, cw, _p(g), _stream()), "loss_tv_bwd")


def loss_inter_bwd(d1, d2, diff, coef, u1, u2, g1, g2):
    check(lib().zt_loss_inter_bwd_f32(_p(d1), _p(d2), _p(diff), d1.shape[1], d1.shape[2], d1.shape[3], coef, _p(u1),
                                      _p(u2), _p(g1), _p(g2), _stream()), "loss_inter_bwd")


def loss_var_bwd(H2, H3, avg1, avg2, coef, t0, t1, t2, g_H2, g_H3):
    check(lib().zt_loss_var_bwd_f32(_p(H2), _p(H3), _p(avg1), _p(avg2), H2.shape[1], H2.shape[2], H2.shape[3], coef,
                                    _p(t0), _p(t1), _p(t2), _p(g_H2), _p(g_H3), _stream()), "loss_var_bwd")


def blur21_bwd(g_in, tmp, g_out):
    _, Cn, H, W = g_in.shape
    check(lib().zt_blur21_bwd_f32(_p(g_in), _p(tmp), _p(g_out), blur_taps(), Cn, H, W, _stream()), "blur21_bwd")


def pair_downsample_bwd(g1, g2, gx):
    _, Cn, H, W = gx.shape
    check(lib().zt_pair_downsample_bwd_f32(_p(g1), _p(g2), Cn, H, W, _p(gx), _stream()), "pair_downsample_bwd")


def div_clamp_bwd(a, add, b, g_out, lo, hi, g_b):
    check(lib().zt_div_clamp_bwd_f32(_p(a), add, _p(b), _p(g_out), a.numel(), lo, hi, _p(g_b), _stream()),
          "div_clamp_bwd")


def mask_pack(gs, outs, c_each, sign, lo, hi, out, cs, HW):
    """out[p][k*c_each+c] = sign * gs[k][c][p] * [lo < outs[k][c][p] < hi]  (outs[k] None: no mask)."""
    g = list(gs) + [None] * (2 - len(gs))
    o = list(outs) + [None] * (2 - len(outs))
    check(lib().zt_mask_pack_nhwc_f32(_p(g[0]), _p(o[0]), _p(g[1]), _p(o[1]), c_each, len(gs), sign, lo, hi, _p(out),
                                      cs, HW, _stream()), "mask_pack")


def unpack_add(g, cs, c_off, c_each, dsts, HW):
    d = list(dsts) + [None] * (2 - len(dsts))
    check(lib().zt_unpack_add_nchw_f32(_p(g), cs, c_off, c_each, len(dsts), _p(d[0]), _p(d[1]), HW, _stream()),
          "unpack_add")


def act_bwd(g, g_cs, y, y_cs, Cn, act, M):
    check(lib().zt_act_bwd_nhwc_f32(_p(g), g_cs, _p(y), y_cs, Cn, ACT[act], M, _stream()), "act_bwd")


def bn_train_bwd(x, x_cs, dout, d_cs, stats, red, gamma, beta, HW, Cn, eps, dx, dx_cs, dgamma, dbeta):
    check(lib().zt_bn_train_bwd_nhwc_f32(_p(x), x_cs, _p(dout), d_cs, _p(stats), _p(red), _p(gamma), _p(beta), HW, Cn,
                                         eps, _p(dx), dx_cs, _p(dgamma), _p(dbeta), _stream()), "bn_train_bwd")


class WGrad:
    """fp32 accumulator of one nn.Conv2d's weight / bias gradient in the kernel layout [kh][kw][cin_p][w_cs]."""

    def __init__(self, wt: ConvW, weight_shape, device):
        self.cout, self.cin, self.kh, self.kw = weight_shape
        self.cin_p, self.w_cs = wt.cin, wt.w_cs
        self.dw = torch.zeros(self.kh, self.kw, self.cin_p, self.w_cs, device=device)
        self.db = torch.zeros(self.w_cs, device=device)

    def zero(self):
        self.dw.zero_()
        self.db.zero_()

    # tensor-core path (csrc/wgrad_tc.cu): 1x1 / 3x3 'same' convolutions with cin, cout <= 64 - every layer of the three
    # Retinex nets.  ZT_WGRAD_TC=0 keeps the fp32 FFMA kernel.
    TC = int(__import__("os").environ.get("ZT_WGRAD_TC", "1"))
    _scratch = {}

    @classmethod
    def _hilo(cls, tag, M, cs, device):
        key = (tag, str(device))
        t = cls._scratch.get(key)
        if t is None or t.numel() < 2 * M * cs:
            t = torch.empty(2 * M * cs, dtype=torch.bfloat16, device=device)
            cls._scratch[key] = t
        return t[:M * cs], t[M * cs:2 * M * cs]

    def accumulate(self, X, x_cs, G, g_cs, H, W, pad):
        if (self.TC and (self.kh, self.kw) in ((1, 1), (3, 3)) and pad == (self.kh // 2, self.kw // 2) and
                self.cin <= 64 and self.cout <= 64):
            M = H * W
            cx, cg = (self.cin + 3) // 4 * 4, (self.cout + 3) // 4 * 4
            xs, gs = (cx + 7) // 8 * 8, (cg + 7) // 8 * 8
            xh, xl = self._hilo("x", M, xs, X.device)
            gh, gl = self._hilo("g", M, gs, X.device)
            check(lib().zt_split_bf16_hilo_colsum(_p(X), x_cs, _p(xh), _p(xl), xs, cx, M, None, _stream()), "split_hilo")
            check(lib().zt_split_bf16_hilo_colsum(_p(G), g_cs, _p(gh), _p(gl), gs, cg, M, _p(self.db), _stream()), "split_hilo")
            check(lib().zt_conv2d_wgrad_tc(_p(xh), _p(xl), xs, _p(gh), _p(gl), gs, _p(self.dw), H, W, self.kh, self.kw,
                                           self.cin, self.cout, self.cin_p, self.w_cs, _stream()), "wgrad_tc")
            return
        check(lib().zt_conv2d_wgrad_nhwc_f32(_p(X), x_cs, _p(G), g_cs, _p(self.dw), _p(self.db), H, W, self.kh, self.kw,
                                             pad[0], pad[1], self.cin_p, self.cout, self.w_cs, _stream()), "wgrad")

    def grads(self):
        """(dweight [cout][cin][kh][kw], dbias [cout]) in torch's parameter layout."""
        return (self.dw[:, :, :self.cin, :self.cout].permute(3, 2, 0, 1).contiguous(), self.db[:self.cout].clone())


def clip_adam(param, grad, exp_avg, exp_avg_sq, sumsq, step, lr, betas, eps, weight_decay, max_norm, grad_scale=1.0):
    check(lib().zt_clip_adam_f32(_p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), param.numel(), _p(sumsq),
                                 grad_scale, max_norm, lr, betas[0], betas[1], eps, weight_decay, step, _stream()),
          "clip_adam")

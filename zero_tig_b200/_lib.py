"""ctypes binding of libzerotig.so (the C ABI declared in include/zerotig.h).

There is deliberately no fallback: if the library is missing or a call fails, a
RuntimeError is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libzerotig.so")

c_fp = C.c_void_p  # device pointers travel as integers


class ZtConv(C.Structure):
    _fields_ = [
        ("in0", c_fp), ("in0_cs", C.c_int), ("c0", C.c_int),
        ("in1", c_fp), ("in1_cs", C.c_int), ("c1", C.c_int),
        ("weight", c_fp), ("w_cs", C.c_int),
        ("bias", c_fp),
        ("out", c_fp), ("out_cs", C.c_int),
        ("B", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Ho", C.c_int), ("Wo", C.c_int), ("cout", C.c_int),
        ("kh", C.c_int), ("kw", C.c_int), ("sh", C.c_int), ("sw", C.c_int), ("ph", C.c_int), ("pw", C.c_int),
        ("act", C.c_int), ("out_scale", C.c_float), ("epi", C.c_int),
        ("aux0", c_fp), ("aux0_cs", C.c_int),
        ("aux1", c_fp), ("aux1_cs", C.c_int),
        ("bplane", c_fp), ("bplane_cs", C.c_int),
        ("acc_scale", C.c_float),
    ]


class ZtConvBf16(C.Structure):
    _fields_ = [
        ("in_", c_fp), ("cin", C.c_int),
        ("weight", c_fp), ("bias", c_fp),
        ("out", c_fp), ("out_cs", C.c_int), ("out_f32", C.c_int),
        ("res", c_fp), ("res_cs", C.c_int),
        ("H", C.c_int), ("W", C.c_int), ("cout", C.c_int), ("kh", C.c_int), ("kw", C.c_int), ("act", C.c_int),
        ("head_w", c_fp), ("head_b", c_fp), ("head_n", C.c_int),
    ]


class ZtRaftUpdate(C.Structure):
    _fields_ = ([("h8", C.c_int), ("w8", C.c_int), ("iters", C.c_int), ("max_ctas", C.c_int)] +
                [(n + s, c_fp) for n in ("corr", "c1", "f1", "c2f", "hm", "rh") for s in ("_hi", "_lo")] +
                [(n, c_fp) for n in ("h", "z", "fhid", "coords1", "flow4", "bp_zr1", "bp_q1", "bp_zr2", "bp_q2")] +
                [("lvl", c_fp * 4), ("lh", C.c_int * 4), ("lw", C.c_int * 4), ("lvl_bf16", C.c_int),
                 ("w", c_fp * 9), ("b", c_fp * 9), ("bn", C.c_int * 9), ("fp16", C.c_int), ("acc_scale", C.c_float * 9)] +
                [(n, c_fp) for n in ("wf1", "bf1", "wfh2", "bfh2", "counter")])


I, L, F, P = C.c_int, C.c_int64, C.c_float, C.c_void_p
IP = C.POINTER(C.c_int)

# name -> argtypes; every function returns int unless listed in _RESTYPE
SIGNATURES = {
    "zt_conv2d_nhwc_f32": [C.POINTER(ZtConv), P],
    "zt_conv2d_tc_nhwc": [C.POINTER(ZtConv), I, I, P],
    "zt_debug_set_tc_trace": [P, I],
    "zt_conv2d_bf16_tma": [C.POINTER(ZtConvBf16), P],
    "zt_pack_nhwc_bf16": [P, P, P, P, I, I, F, P, I, L, P],
    "zt_space_to_depth2_f32": [P, I, I, I, I, I, P, P],
    "zt_raft_convf1_f32": [P, I, P, P, P, I, I, I, P],
    "zt_conv3x3_n2_f32": [P, I, I, P, P, P, I, I, I, P],
    "zt_flow_head2_update_f32": [P, I, I, P, P, P, P, I, P, I, I, I, P],
    "zt_raft_update": [C.POINTER(ZtRaftUpdate), P],
    "zt_debug_set_raft_update_trace": [P, I],
    "zt_split_bf16_hilo": [P, I, P, P, I, I, L, P],
    "zt_pack_tc_rows_f32": [P, I, I, I, I, P, P],
    "zt_pack_tc_weights_f32": [P, I, I, I, I, I, I, I, I, F, P, P],
    "zt_instnorm_nhwc_f32": [P, I, P, I, P, I, P, I, I, I, F, I, I, P],
    "zt_instnorm_nhwc_f32_noalloc": [P, I, P, I, P, I, I, I, I, F, I, I, P],
    "zt_transpose_f32": [P, I, I, I, P, I, P],
    "zt_avgpool2_f32": [P, P, L, I, I, P],
    "zt_corr_lookup_f32": [P, P, P, P, IP, IP, P, P, I, I, P],
    "zt_corr_lookup_bf16": [P, P, P, P, IP, IP, P, P, I, I, P],
    "zt_corr_pyramid_bf16": [P, I, I, I, P, P, P, P, P, L, P],
    "zt_corr_pyramid_scratch_bytes": [I, I],
    "zt_corr_pyramid_f32": [P, I, I, I, P, P, P, P, P, L, P],
    "zt_convex_upsample_f32": [P, I, P, I, P, I, I, P],
    "zt_flow_update_f32": [P, P, I, P, I, P, I, I, I, P],
    "zt_resize255_f32": [P, I, I, I, I, I, P, I, I, P, P, P],
    "zt_equalize_lut": [P, P, P],
    "zt_equalize_apply_pack": [P, P, I, I, I, I, P, P, P],
    "zt_warp_pair_f32": [P, I, I, P, P, P, P, I, I, I, P],
    "zt_pack_nhwc_f32": [P, P, P, P, I, I, F, P, I, L, P],
    "zt_retinex_l2_f32": [P, P, I, P, L, F, P],
    "zt_retinex_h2_f32": [P, P, I, P, P, L, F, P],
    "zt_retinex_h5_f32": [P, P, P, I, P, P, L, F, P],
    "zt_quantize_u8_hwc": [P, P, I, L, P],
    "zt_u8_hwc_to_f32": [P, P, I, L, P],
    "zt_bn_train_nhwc_f32": [P, I, P, I, P, I, P, P, P, P, P, I, I, F, F, P],
    "zt_pair_downsample_f32": [P, P, P, I, I, I, F, P],
    "zt_sub_clamp_f32": [P, F, P, I, P, I, L, F, F, I, P],
    "zt_div_clamp_f32": [P, F, P, P, L, F, F, P],
    "zt_texture_diff_f32": [P, P, P, I, I, P],
    "zt_blur21_f32": [P, P, P, C.POINTER(C.c_float), I, I, I, P],
    "zt_loss_illum_f32": [P, P, I, L, P, P, P],
    "zt_loss_smooth_f32": [P, P, P, I, I, P, P],
    "zt_loss_tv_f32": [P, I, I, I, P, P],
    "zt_mse_sum_f32": [P, P, F, L, P, P],
    "zt_loss_inter_f32": [P, P, P, I, I, I, P, P],
    "zt_loss_var_f32": [P, P, P, P, I, I, I, P, P],
    "zt_mse_bwd_f32": [P, P, F, L, F, P, P, P],
    "zt_loss_illum_bwd_f32": [P, P, I, L, P, F, F, P, P],
    "zt_loss_smooth_bwd_f32": [P, P, I, I, C.POINTER(C.c_float), P, P],
    "zt_loss_tv_bwd_f32": [P, I, I, I, F, F, P, P],
    "zt_loss_inter_bwd_f32": [P, P, P, I, I, I, F, P, P, P, P, P],
    "zt_loss_var_bwd_f32": [P, P, P, P, I, I, I, F, P, P, P, P, P, P],
    "zt_blur21_bwd_f32": [P, P, P, C.POINTER(C.c_float), I, I, I, P],
    "zt_pair_downsample_bwd_f32": [P, P, I, I, I, P, P],
    "zt_div_clamp_bwd_f32": [P, F, P, P, L, F, F, P, P],
    "zt_mask_pack_nhwc_f32": [P, P, P, P, I, I, F, F, F, P, I, L, P],
    "zt_unpack_add_nchw_f32": [P, I, I, I, I, P, P, L, P],
    "zt_act_bwd_nhwc_f32": [P, I, P, I, I, I, L, P],
    "zt_bn_train_bwd_nhwc_f32": [P, I, P, I, P, P, P, P, I, I, F, P, I, P, P, P],
    "zt_conv2d_wgrad_nhwc_f32": [P, I, P, I, P, P, I, I, I, I, I, I, I, I, I, P],
    "zt_split_bf16_hilo_colsum": [P, I, P, P, I, I, L, P, P],
    "zt_conv2d_wgrad_tc": [P, P, I, P, P, I, P, I, I, I, I, I, I, I, I, P],
    "zt_clip_adam_f32": [P, P, P, P, L, P, F, F, F, F, F, F, F, I, P],
    "zt_last_error": [],
    "zt_version": [],
    "zt_launch_count": [],
    "zt_add_launches": [L],
}
_RESTYPE = {"zt_last_error": C.c_char_p, "zt_launch_count": C.c_int64, "zt_corr_pyramid_scratch_bytes": C.c_int64}

_lib = None


def lib():
    """Load (once) and return the ctypes handle; raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libzerotig.so is not built (%s). Run `python -m zero_tig_b200.build` "
                "(needs nvcc); there is no CPU fallback." % LIB_PATH)
        h = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(h, name)  # AttributeError if the symbol is missing
            fn.argtypes = args
            fn.restype = _RESTYPE.get(name, C.c_int)
        _lib = h
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().zt_last_error()
        raise RuntimeError("libzerotig %s failed (code %d): %s" % (what, rc, msg.decode() if msg else "?"))


def launch_count():
    return int(lib().zt_launch_count())

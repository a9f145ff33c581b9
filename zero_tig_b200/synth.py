"""Deterministic synthetic weights and clips (the shipped checkpoints and datasets are not
available offline - SURVEY F3).  Used by tests, bench.py and the golden-vector generator so
that every party sees the same tensors without storing them.
"""
import argparse
import math

import numpy as np
import torch


def make_args(dataset="RLV", of_scale=3, model_pretrain=None):
    """argparse.Namespace like the reference CLIs pass (predict.py:23-38, train.py:15-29)."""
    return argparse.Namespace(dataset=dataset, of_scale=of_scale, model_pretrain=model_pretrain)


def synthetic_state_dict(template: dict, seed: int = 1234) -> dict:
    """Fill every tensor of ``template`` (a state_dict, e.g. the reference Network's)
    deterministically from its key: conv weights ~ N(0, sqrt(2/fan_out)) for RAFT and
    N(0, 0.02) for the Retinex nets (train.py:82-84 init law), small non-zero biases and
    non-trivial BatchNorm affine / running statistics so folding is exercised."""
    out = {}
    for i, (k, v) in enumerate(template.items()):
        g = torch.Generator().manual_seed(seed * 100003 + i)
        shape = tuple(v.shape)
        if k.endswith("num_batches_tracked"):
            t = torch.tensor(7, dtype=v.dtype)
        elif k.endswith("running_mean"):
            t = 0.1 * torch.randn(shape, generator=g)
        elif k.endswith("running_var"):
            t = 0.8 + 0.4 * torch.rand(shape, generator=g)
        elif v.dim() == 4:
            if k.startswith("raft.update_block."):   # torch default conv init: U(+-1/sqrt(fan_in))
                bound = 1.0 / math.sqrt(shape[1] * shape[2] * shape[3])
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif k.startswith("raft."):              # extractor.py:149-156 kaiming fan_out
                fan_out = shape[0] * shape[2] * shape[3]
                t = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_out)
            else:
                t = torch.randn(shape, generator=g) * 0.02
        elif k.endswith(".weight"):            # norm scale
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        else:                                   # biases
            t = 0.02 * torch.randn(shape, generator=g)
        out[k] = t.to(v.dtype)
    # aliased tensors (enhance.blocks.N.* == enhance.conv.*, downsample.1 == norm3) must agree
    for k in list(out):
        if k.startswith("enhance.blocks."):
            out[k] = out["enhance.conv." + k.split(".", 3)[3]]
        if ".downsample.1." in k:
            out[k] = out[k.replace(".downsample.1.", ".norm3.")]
    return out


def synthetic_clip(n_frames: int, H: int, W: int, seed: int = 0, underwater: bool = False) -> torch.Tensor:
    """Low-light clip [n,3,H,W] fp32 on the k/255 grid: band-limited texture + rectangles,
    translated by (2t, t) px with 0.5 %/frame zoom, x0.15 darkening, Gaussian + Poisson noise
    (SURVEY 8d)."""
    rng = np.random.default_rng(seed)
    pad = 4 * n_frames + 16
    Hb, Wb = H + 2 * pad, W + 2 * pad
    # band-limited noise: upsampled coarse noise (sigma ~ 8 px)
    ch, cw = Hb // 8 + 2, Wb // 8 + 2
    coarse = torch.from_numpy(rng.random((1, 3, ch, cw)).astype(np.float32))
    base = torch.nn.functional.interpolate(coarse, size=(Hb, Wb), mode="bicubic", align_corners=False).clamp(0, 1)
    for _ in range(6):
        y0, x0 = int(rng.integers(0, Hb - 40)), int(rng.integers(0, Wb - 40))
        hh, ww = int(rng.integers(16, 40 + Hb // 6)), int(rng.integers(16, 40 + Wb // 6))
        base[0, :, y0:y0 + hh, x0:x0 + ww] = torch.from_numpy(rng.random(3).astype(np.float32)).view(3, 1, 1)
    frames = []
    ys = torch.linspace(-1, 1, H).view(H, 1).expand(H, W)
    xs = torch.linspace(-1, 1, W).view(1, W).expand(H, W)
    for t in range(n_frames):
        zoom = 1.0 + 0.005 * t
        gx = (xs / zoom) * (W / Wb) + (2.0 * t) * 2.0 / Wb
        gy = (ys / zoom) * (H / Hb) + (1.0 * t) * 2.0 / Hb
        grid = torch.stack([gx, gy], -1)[None]
        f = torch.nn.functional.grid_sample(base, grid, mode="bilinear", padding_mode="border", align_corners=False)
        if underwater:
            f = f * torch.tensor([0.25, 0.7, 1.0]).view(1, 3, 1, 1)
        f = f * 0.15
        lam = (f * 255.0 * 4.0).numpy()
        noisy = rng.poisson(lam).astype(np.float32) / 4.0 + rng.normal(0, 3.0, lam.shape).astype(np.float32)
        q = np.clip(np.round(noisy), 0, 255).astype(np.uint8)
        frames.append(torch.from_numpy(q.astype(np.float32) / 255.0))
    return torch.cat(frames, 0)


def synthetic_clip_mt(n_frames: int, H: int, W: int, seed: int = 0, underwater: bool = False, start: int = 0,
                      threads: int = 0, as_u8: bool = False) -> torch.Tensor:
    """The same image model as ``synthetic_clip`` (translating / zooming band-limited texture, x0.15 darkening,
    Poisson + Gaussian noise, values on the k/255 grid) but with one independent noise generator per frame, so the
    frames [start, start + n_frames) of a long clip can be made on several host threads and a rank can make only
    its own frame range.  For benchmarks (64-frame 1080p / 4K clips); the tests and goldens keep ``synthetic_clip``.
    ``as_u8`` returns the decoded 8-bit HWC frames [n,H,W,3] (what PIL hands to ``ToTensor``) instead of fp32 NCHW."""
    import concurrent.futures as cf
    import os
    total = start + n_frames
    rng = np.random.default_rng(seed)
    pad = 4 * 64 + 16                # independent of the range requested: every caller sees the same clip
    Hb, Wb = H + 2 * pad, W + 2 * pad
    ch, cw = Hb // 8 + 2, Wb // 8 + 2
    coarse = torch.from_numpy(rng.random((1, 3, ch, cw)).astype(np.float32))
    base = torch.nn.functional.interpolate(coarse, size=(Hb, Wb), mode="bicubic", align_corners=False).clamp(0, 1)
    for _ in range(6):
        y0, x0 = int(rng.integers(0, Hb - 40)), int(rng.integers(0, Wb - 40))
        hh, ww = int(rng.integers(16, 40 + Hb // 6)), int(rng.integers(16, 40 + Wb // 6))
        base[0, :, y0:y0 + hh, x0:x0 + ww] = torch.from_numpy(rng.random(3).astype(np.float32)).view(3, 1, 1)
    ys = torch.linspace(-1, 1, H).view(H, 1).expand(H, W)
    xs = torch.linspace(-1, 1, W).view(1, W).expand(H, W)
    gains = torch.tensor([0.25, 0.7, 1.0]).view(1, 3, 1, 1)

    def make(t):
        tt = t % 64                       # the motion pattern repeats every 64 frames (bounded padding)
        zoom = 1.0 + 0.005 * tt
        gx = (xs / zoom) * (W / Wb) + (2.0 * tt) * 2.0 / Wb
        gy = (ys / zoom) * (H / Hb) + (1.0 * tt) * 2.0 / Hb
        f = torch.nn.functional.grid_sample(base, torch.stack([gx, gy], -1)[None], mode="bilinear",
                                            padding_mode="border", align_corners=False)
        if underwater:
            f = f * gains
        lam = (f * (0.15 * 255.0 * 4.0)).numpy()
        r = np.random.default_rng([seed, t])
        noisy = r.poisson(lam).astype(np.float32) / 4.0 + r.normal(0, 3.0, lam.shape).astype(np.float32)
        return np.clip(np.round(noisy), 0, 255).astype(np.uint8)[0]          # [3,H,W] uint8

    nthr = threads or min(16, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(max_workers=nthr) as ex:
        frames = list(ex.map(make, range(start, total)))
    q = torch.from_numpy(np.stack(frames, 0))
    if as_u8:
        return q.permute(0, 2, 3, 1).contiguous()
    return q.float().div(255)

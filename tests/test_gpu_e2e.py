"""GPU: the whole drop-in path (Finetunemodel.forward, RAFT.forward, update_cache) against
the reference golden clips and the CPU oracle.  Tolerances are BASELINE.json's: flow EPE
<= 0.01 px, frames >= 50 dB PSNR."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, epe, psnr

pytestmark = pytest.mark.gpu

EPE_TOL = 0.01     # px, BASELINE.json north_star
PSNR_TOL = 50.0    # dB


@pytest.fixture(scope="module")
def model_factory(weights, tmp_path_factory):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    ck = tmp_path_factory.mktemp("w") / "synthetic-RLV.pt"
    torch.save(weights, ck)

    def make(of_scale):
        m = Finetunemodel(synth.make_args(of_scale=of_scale, model_pretrain=str(ck))).cuda().eval()
        m.raft.load_state_dict({k[5:]: v for k, v in weights.items() if k.startswith("raft.")})
        return m
    return make


@pytest.mark.parametrize("tag", ["s1_128x160", "s3_390x400"])
def test_finetune_clip_vs_reference_golden(tag, model_factory):
    from zero_tig_b200 import synth, _lib
    g = np.load(os.path.join(GOLDEN, f"infer_{tag}.npz"))
    H, W, of_scale, n, stride = [int(v) for v in g["meta"][:5]]
    clip = synth.synthetic_clip(n, H, W, seed=3).cuda()
    m = model_factory(of_scale)
    before = _lib.launch_count()
    for t in range(n):
        m.is_new_seq = (t == 0)
        H2, H3, s3 = m(clip[t:t + 1])
        assert H2.shape == (1, 3, H, W) and H3.is_cuda
        for name, v in (("H2", H2), ("H3", H3), ("s3", s3)):
            ref = torch.from_numpy(g[f"{name}_{t}"])
            p = psnr(v[0, :, ::stride, ::stride].cpu(), ref)
            assert p >= PSNR_TOL, (name, t, p)
    assert _lib.launch_count() - before > 100, "kernels did not launch through libzerotig"
    eng = m.raft.engine()
    # flow of the last pair: padded-size flow_up, strided like the fixture
    fu = [v for k, v in eng.ws.bufs.items() if k[0] == "flow_up"][0]
    e = epe(fu[0, :, ::stride, ::stride].cpu(), torch.from_numpy(g["flow_up_last"]))
    assert e <= EPE_TOL, e
    fl = eng.last["flow4"][:, :2].reshape(eng.last["h8"], eng.last["w8"], 2).permute(2, 0, 1).cpu()
    assert epe(fl, torch.from_numpy(g["flow_low_last"])) <= EPE_TOL / 8


def test_raft_forward_api_vs_oracle(weights, model_factory):
    """RAFT.forward(image1, image2) public signature; trace intermediates vs the oracle."""
    import zt_oracle as O
    m = model_factory(1)
    g = torch.Generator().manual_seed(4)
    base = torch.rand(1, 3, 150, 170, generator=g)
    im1 = (base * 255).round()
    im2 = (torch.roll(base, (2, 3), (2, 3)) * 255).round()
    tr = {}
    fl_ref, fu_ref = O.raft_forward(im1, im2, weights, trace=tr)
    fl, fu = m.raft(im1.cuda(), im2.cuda(), iters=12, test_mode=True)
    assert fu.shape == fu_ref.shape == (1, 2, 152, 176)        # padded size is returned (Q3)
    last = m.raft.engine().last
    N = last["N"]
    fm = last["fmap"].reshape(2, last["h8"], last["w8"], 256).permute(0, 3, 1, 2).cpu()
    assert (fm[0] - tr["fmap1"][0]).abs().max().item() < 2e-4
    assert (fm[1] - tr["fmap2"][0]).abs().max().item() < 2e-4
    for i in range(4):
        ref = tr["pyr"][i].reshape(N, -1)
        # correlation values are O(20): 1e-4 of the level's range (bf16 hi/lo split convs upstream)
        assert (last["levels"][i].cpu() - ref).abs().max().item() < 1e-4 * ref.abs().max().item()
    assert epe(fu.cpu(), fu_ref) <= EPE_TOL
    assert epe(fl.cpu(), fl_ref) <= EPE_TOL / 8


def test_update_cache_matches_oracle(weights, model_factory):
    import zt_oracle as O
    m = model_factory(3)
    g = torch.Generator().manual_seed(8)
    H, W = 384, 402
    prev = torch.rand(1, 3, H, W, generator=g)
    prev_s = torch.rand(1, 3, H, W, generator=g)
    L2 = torch.clamp(torch.roll(prev, (3, 6), (2, 3)) * 0.2, 1e-4, 1)
    ra, rb = O.update_cache(prev, prev_s, L2, weights, 3)
    a, b = m.update_cache(prev.cuda(), prev_s.cuda(), L2.cuda())
    assert psnr(a.cpu(), ra) >= PSNR_TOL and psnr(b.cpu(), rb) >= PSNR_TOL


def test_new_sequence_resets_state(model_factory):
    """is_new_seq=True must not read the previous state (model.py:319-321,330-332)."""
    from zero_tig_b200 import synth
    m = model_factory(3)
    clip = synth.synthetic_clip(2, 384, 400, seed=5).cuda()
    m.is_new_seq = True
    a = [t.clone() for t in m(clip[0:1])]
    m.is_new_seq = False
    m(clip[1:2])
    m.is_new_seq = True
    b = m(clip[0:1])
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_too_small_frame_is_refused(model_factory):
    """Reference yields NaN when the coarsest pyramid level is 1 px; we raise."""
    m = model_factory(3)
    x = torch.rand(1, 3, 256, 256).cuda()
    m.is_new_seq = True
    m(x)
    m.is_new_seq = False
    with pytest.raises(RuntimeError, match="too small|NaN"):
        m(x)


@pytest.mark.parametrize("u8,in_u8,graph", [(False, False, True), (True, False, True), (True, True, True),
                                             (False, False, False)])
def test_frame_pipeline_matches_direct_calls(model_factory, u8, in_u8, graph):
    """FramePipeline (pinned host frames in, host H2/H3 out, copies on side streams, steady-state frames replayed as
    one CUDA graph) returns, in order, exactly what per-frame eager model calls return; `quantize_u8` equals
    predict.py:57-61's clip(x*255).astype(uint8); `input_u8` (8-bit HWC upload + ToTensor on the device) feeds the
    model bit-identical frames."""
    from zero_tig_b200 import synth
    from zero_tig_b200.pipeline import FramePipeline
    H, W, n = 128, 160, 7
    clip = synth.synthetic_clip(n, H, W, seed=11)
    a, b = model_factory(1), model_factory(1)
    want = []
    for t in range(n):
        a.is_new_seq = (t == 0)
        H2, H3, _ = a(clip[t:t + 1].cuda())
        want.append((H2.cpu(), H3.cpu()))
    got = {}
    pipe = FramePipeline(b, H, W, depth=2, quantize_u8=u8, input_u8=in_u8, graph=graph)
    if in_u8:
        frames = [(clip[t].permute(1, 2, 0) * 255.0).round().to(torch.uint8).contiguous().pin_memory() for t in range(n)]
        assert torch.equal(frames[3].float().div(255).permute(2, 0, 1), clip[3])      # ToTensor round trip is exact
    else:
        frames = [clip[t:t + 1].pin_memory() for t in range(n)]
    pipe.run(frames, lambda i, h2, h3: got.__setitem__(i, (h2.clone(), h3.clone())))
    assert sorted(got) == list(range(n))
    for t in range(n):
        for g_, w_ in zip(got[t], want[t]):
            if u8:
                ref = np.clip(w_[0].numpy().transpose(1, 2, 0) * 255.0, 0, 255.0).astype("uint8")
                assert g_.shape == (H, W, 3) and np.array_equal(g_.numpy(), ref), t
            else:
                assert torch.equal(g_, w_), t
    with pytest.raises(ValueError):
        pipe.run([torch.zeros(2, 2)], lambda *a_: None)


def test_u8_converters_bit_exact():
    """zt_u8_hwc_to_f32 == transforms.ToTensor (value / 255, true division) and zt_quantize_u8_hwc ==
    clip(x*255, 0, 255).astype(uint8) (predict.py:57-61), vector and scalar paths."""
    from zero_tig_b200 import ops
    g = torch.Generator().manual_seed(2)
    for H, W in ((16, 16), (5, 7)):                 # HW % 4 == 0 -> x4 kernels, else scalar kernels
        img = torch.randint(0, 256, (H, W, 3), generator=g, dtype=torch.uint8)
        if H == 16:
            img.view(-1)[:256] = torch.arange(256, dtype=torch.uint8)
        out = torch.empty(1, 3, H, W, device="cuda")
        ops.u8_hwc_to_f32(img.cuda(), out, 3, H * W)
        assert torch.equal(out.cpu()[0], img.float().div(255).permute(2, 0, 1))
        x = torch.rand(1, 3, H, W, generator=g) * 1.2 - 0.1
        x.view(-1)[:4] = torch.tensor([0.0, 1.0, 254.999 / 255, 1.0 / 255])
        q = torch.empty(H, W, 3, dtype=torch.uint8, device="cuda")
        ops.quantize_u8_hwc(x.cuda(), q, 3, H * W)
        ref = np.clip(x[0].numpy().transpose(1, 2, 0) * 255.0, 0, 255.0).astype("uint8")
        assert np.array_equal(q.cpu().numpy(), ref)


@pytest.mark.parametrize("H,W,dataset,n", [(1080, 1920, "RLV", 3), (2160, 3840, "underwater", 2)])
def test_full_size_clip_vs_oracle(H, W, dataset, n, weights, tmp_path):
    """BASELINE.json configs[1] (1080p RLV) and configs[3] (4K underwater, BVI-Coral-style `is_WB` weights) at their
    full frame sizes, of_scale 3, against the CPU oracle on the same synthetic clip: every returned frame >= 50 dB,
    last flow_up EPE <= 0.01 px.  The oracle needs ~2 s (1080p) / ~8 s (4K) per frame on the box's host cores."""
    import zt_oracle as O
    from zero_tig_b200 import synth, _lib
    from zero_tig_b200.model.model import Finetunemodel
    ck = tmp_path / "synthetic.pt"
    torch.save(weights, ck)
    m = Finetunemodel(synth.make_args(dataset=dataset, of_scale=3, model_pretrain=str(ck))).cuda().eval()
    m.raft.load_state_dict({k[5:]: v for k, v in weights.items() if k.startswith("raft.")})
    clip = synth.synthetic_clip(n, H, W, seed=3, underwater=(dataset == "underwater"))
    st, tr = O.InferState(), {}
    before = _lib.launch_count()
    for t in range(n):
        m.is_new_seq = (t == 0)
        got = m(clip[t:t + 1].cuda())
        ref = O.finetune_forward(clip[t:t + 1], weights, st, t == 0, 3, trace=tr)
        for name, v, r in zip(("H2", "H3", "s3"), got, ref):
            assert v.shape == (1, 3, H, W)
            p = psnr(v.cpu(), r)
            assert p >= PSNR_TOL, (name, t, p)
    assert _lib.launch_count() - before > 100
    fu = [v for k, v in m.raft.engine().ws.bufs.items() if k[0] == "flow_up"][0]
    assert fu.shape == tr["flow_up"].shape
    assert epe(fu.cpu(), tr["flow_up"]) <= EPE_TOL


def test_multi_shard_pipeline_equals_per_shard_runs(model_factory):
    """Two frame-range shards interleaved on two streams of one GPU (MultiShardPipeline) return, per shard and in
    order, bit-identical frames to a lone FramePipeline over the same (halo + range) sub-clip."""
    from zero_tig_b200 import synth
    from zero_tig_b200.pipeline import FramePipeline, MultiShardPipeline
    H, W = 128, 160
    clip = synth.synthetic_clip(11, H, W, seed=13)
    # shard 0: frames 0..5; shard 1: halo frame 5 + frames 6..10 (shard.plan's cut)
    subs = [[clip[t:t + 1].pin_memory() for t in range(0, 6)], [clip[t:t + 1].pin_memory() for t in range(5, 11)]]
    want = []
    for frames in subs:
        got = []
        FramePipeline(model_factory(1), H, W).run(frames, lambda i, a, b: got.append((i, a.clone(), b.clone())))
        want.append(got)
    multi = [[], []]
    mp = MultiShardPipeline([model_factory(1), model_factory(1)], H, W)
    mp.run(subs, lambda s, i, a, b: multi[s].append((i, a.clone(), b.clone())))
    torch.cuda.synchronize()
    for s in range(2):
        assert [i for i, _, _ in multi[s]] == list(range(6))
        for (i, a, b), (j, ra, rb) in zip(multi[s], want[s]):
            assert i == j and torch.equal(a, ra) and torch.equal(b, rb), (s, i)

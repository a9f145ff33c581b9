"""CPU: the C-ABI library loads and exports every symbol include/zerotig.h declares; the
host mirrors keep the reference interface; compute refuses to run without CUDA."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "zerotig.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(zt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from zero_tig_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "libzerotig.so not built: run __graft_entry__.build()"
    h = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(h, n), n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert _lib.lib().zt_version() >= 100


def test_arg_validation_without_gpu():
    """Argument errors are reported through the int code + zt_last_error (no launch needed)."""
    from zero_tig_b200 import _lib
    L = _lib.lib()
    rc = L.zt_warp_pair_f32(None, 0, 0, None, None, None, None, 3, 8, 8, None)
    assert rc == 1 and b"warp_pair" in L.zt_last_error()
    with pytest.raises(RuntimeError):
        _lib.check(rc, "warp")


def test_module_interface_matches_reference(weights, tmp_path):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel, Network
    ck = tmp_path / "w.pt"
    torch.save(weights, ck)
    m = Finetunemodel(synth.make_args(model_pretrain=str(ck)))
    assert m.of_scale == 3 and m.is_new_seq is True and m.last_H3 is None
    # raft.* keys are dropped by the key-filtered merge (constructed after loading), others load
    assert torch.equal(m.enhance.in_conv[0].weight, weights["enhance.in_conv.0.weight"])
    assert m.enhance.blocks[2] is m.enhance.conv
    for attr in ("forward", "update_H3", "update_cache", "load_raft"):
        assert callable(getattr(m, attr))
    n = Network(synth.make_args(dataset="underwater"))
    assert n.is_WB and not any(p.requires_grad for p in n.raft.parameters())
    assert sum(p.numel() for p in n.parameters() if p.requires_grad) == 92620
    assert sum(p.numel() for p in n.raft.parameters()) == 5257536


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only check")
def test_no_cpu_fallback(weights, tmp_path):
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    ck = tmp_path / "w.pt"
    torch.save(weights, ck)
    m = Finetunemodel(synth.make_args(model_pretrain=str(ck)))
    # train-mode BatchNorm is not what the inference engine computes (it folds the running statistics): refuse
    with pytest.raises(RuntimeError, match="eval"):
        m(torch.rand(1, 3, 64, 64))
    m.eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.rand(1, 3, 64, 64))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the loud failure on a box without a GPU")
def test_pipelines_refuse_to_run_without_gpu(weights, tmp_path):
    """FramePipeline / MultiShardPipeline only move pinned buffers around the CUDA path: no device, no pipeline."""
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    from zero_tig_b200.pipeline import FramePipeline, MultiShardPipeline
    ck = tmp_path / "w.pt"
    torch.save(weights, ck)
    m = Finetunemodel(synth.make_args(model_pretrain=str(ck)))
    with pytest.raises(RuntimeError, match="CUDA"):
        FramePipeline(m, 64, 64)
    with pytest.raises(RuntimeError, match="CUDA"):
        MultiShardPipeline([m, m], 64, 64)
    with pytest.raises(ValueError):
        MultiShardPipeline([], 64, 64)

import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# torch references computed on the GPU must be true fp32: no TF32 in cuDNN / cuBLAS
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def weights():
    """Synthetic weights (same generator the golden script used), keyed like the reference."""
    import json
    from zero_tig_b200 import synth
    with open(os.path.join(GOLDEN, "state_dict_keys.json")) as f:
        keys = json.load(f)
    # insertion order of the reference state_dict matters for the per-key seeds: rebuild it
    from zero_tig_b200.model.model import Network
    tmpl = Network(synth.make_args()).state_dict()
    assert {k: list(v.shape) for k, v in tmpl.items()} == keys
    return synth.synthetic_state_dict(tmpl, seed=1234)


def psnr(a, b):
    mse = torch.mean((a.double() - b.double()) ** 2).item()
    return 200.0 if mse == 0 else 10.0 * torch.log10(torch.tensor(1.0 / mse)).item()


def epe(a, b):
    return torch.sqrt(((a.double() - b.double()) ** 2).sum(-3)).mean().item()

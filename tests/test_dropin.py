"""The drop-in boundary as the reference CLIs use it: ``from model.model import Finetunemodel``
(predict.py:10) / ``from model.model import *`` (train.py:12) must bind to THIS repository's classes even though
the script's own directory - which holds the reference's ``model/`` package - is ``sys.path[0]``.

CPU tests: import resolution from a foreign working directory, through ``python -m zero_tig_b200.run``, and (when
/root/reference is present, i.e. in the build container) the UNMODIFIED predict.py driven up to the first model
call, which must fail loudly with this package's "CUDA only" error - proof that the real CLI, the real
dataloader and the new model are wired together.  GPU test: a predict.py-like caller (tests/callers/) end to end."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT

REF = "/root/reference"
CALLERS = os.path.join(ROOT, "tests", "callers")
STUBS = os.path.join(ROOT, "tests", "stubs")


def _env(extra_path=()):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([ROOT, *extra_path, env.get("PYTHONPATH", "")])
    return env


def _make_rlv_tree(root, n, H=96, W=128):
    """input/S01/low_light_10/0000N.png + test_list.txt (multi_read_data.py:91-125 layout)."""
    from PIL import Image
    from zero_tig_b200 import synth
    d = os.path.join(root, "input", "S01", "low_light_10")
    os.makedirs(d, exist_ok=True)
    clip = synth.synthetic_clip(n, H, W, seed=21)
    for t in range(n):
        im = (clip[t].permute(1, 2, 0).numpy() * 255.0 + 0.5).astype(np.uint8)
        Image.fromarray(im).save(os.path.join(d, "%05d.png" % (t + 1)))
    for name in ("test_list.txt", "train_list.txt"):
        with open(os.path.join(root, name), "w") as f:
            f.write("S01\n")
    return clip


def _checkpoint(path, weights):
    torch.save({k: v for k, v in weights.items() if not k.startswith("raft.")}, path)


def test_install_puts_dropin_first_from_foreign_cwd(tmp_path):
    """`model.model` resolves to zero_tig_b200/dropin even when the CWD (sys.path[0] == '') holds another `model`
    package - the situation of running inside the reference checkout."""
    cwd = REF if os.path.isdir(os.path.join(REF, "model")) else CALLERS
    code = ("import zero_tig_b200.run as r; r.install(%r); import model.model as m, model.RAFT.raft as rr; "
            "print(m.__file__); print(m.Finetunemodel.__module__, m.Network.__module__); print(rr.RAFT.__module__);"
            "assert hasattr(m, 'nn') and hasattr(m, 'np') and hasattr(m, 'F')") % cwd
    r = subprocess.run([sys.executable, "-c", code], cwd=cwd, env=_env(), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert os.path.samefile(lines[0], os.path.join(ROOT, "zero_tig_b200", "dropin", "model", "model.py"))
    assert lines[1] == "zero_tig_b200.model.model zero_tig_b200.model.model"
    assert lines[2] == "zero_tig_b200.model.RAFT.raft"


def test_plain_script_run_is_shadowed_by_script_dir():
    """Documents WHY the launcher exists: `PYTHONPATH=<dropin> python script.py` loses to the script directory."""
    dropin = os.path.join(ROOT, "zero_tig_b200", "dropin")
    r = subprocess.run([sys.executable, os.path.join(CALLERS, "predict_like.py"), "--help"], env=_env([dropin]),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "shadowed the zero_tig_b200 drop-in" in r.stderr


def test_launcher_wins_over_script_dir(tmp_path):
    """`python -m zero_tig_b200.run script.py --help`: the script's `from model.model import Finetunemodel` binds to
    the drop-in (argparse then prints the SCRIPT's usage and exits 0)."""
    r = subprocess.run([sys.executable, "-m", "zero_tig_b200.run", os.path.join(CALLERS, "predict_like.py"), "--help"],
                       cwd=str(tmp_path), env=_env(), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "predict-like" in r.stdout and "--lowlight_images_path" in r.stdout


@pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "predict.py")), reason="reference tree not present")
def test_unmodified_predict_py_reaches_the_b200_model(tmp_path, weights):
    """The UNMODIFIED /root/reference/predict.py under the launcher, CWD = the reference checkout: its argparse, its
    dataloader (PIL resize to 1080p) and its loop run; the first `model(input)` reaches zero_tig_b200's engine.
    Without a GPU that call must raise this package's own error (no CPU fallback); with one it must finish."""
    _make_rlv_tree(str(tmp_path / "data"), 2)
    ck = tmp_path / "synthetic-RLV.pt"
    _checkpoint(ck, weights)
    r = subprocess.run([sys.executable, "-m", "zero_tig_b200.run", os.path.join(REF, "predict.py"),
                        "--lowlight_images_path", str(tmp_path / "data"), "--save", str(tmp_path / "out"),
                        "--model_pretrain", str(ck)], cwd=REF, env=_env([STUBS]), capture_output=True, text=True,
                       timeout=900)
    assert "Total image number:  2" in r.stdout, r.stdout + r.stderr
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
        assert sorted(os.listdir(tmp_path / "out" / "S01" / "low_light_10")) == [
            "00001_denoise.png", "00001_enhance.png", "00002_denoise.png", "00002_enhance.png"]
    else:
        assert r.returncode != 0
        assert "zero_tig_b200 runs on CUDA only" in r.stderr, r.stderr
        assert "zero_tig_b200/engine.py" in r.stderr


@pytest.mark.gpu
def test_predict_like_cli_end_to_end(tmp_path, weights):
    """predict.py's calling sequence (tests/callers/predict_like.py: default device = cuda before construction,
    DataLoader tuples, is_new_seq rule, two PNGs per frame) through the launcher; the PNGs must equal what direct
    module calls on the same frames give after predict.py:57-61's uint8 conversion."""
    from PIL import Image
    from zero_tig_b200 import synth
    from zero_tig_b200.model.model import Finetunemodel
    n, H, W = 3, 390, 400
    clip = _make_rlv_tree(str(tmp_path / "data"), n, H, W)
    ck = tmp_path / "synthetic-RLV.pt"
    _checkpoint(ck, weights)
    r = subprocess.run([sys.executable, "-m", "zero_tig_b200.run", os.path.join(CALLERS, "predict_like.py"),
                        "--lowlight_images_path", str(tmp_path / "data"), "--save", str(tmp_path / "out"),
                        "--model_pretrain", str(ck), "--size", "%dx%d" % (W, H)], cwd=str(tmp_path), env=_env(),
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr
    mod = [ln.split("=", 1)[1] for ln in r.stdout.splitlines() if ln.startswith("MODEL_MODULE_FILE=")][0]
    assert os.path.samefile(mod, os.path.join(ROOT, "zero_tig_b200", "dropin", "model", "model.py"))
    out = tmp_path / "out" / "S01" / "low_light_10"
    assert sorted(os.listdir(out)) == sorted("%05d_%s.png" % (t + 1, k) for t in range(n) for k in ("denoise", "enhance"))
    # same frames (as the PNG round trip delivers them) through direct calls; Finetunemodel builds a random RAFT
    #  - seed both the same way the CLI process did not, so compare frame 0 (no flow) bit-exactly and
    # check the others are valid images
    m = Finetunemodel(synth.make_args(of_scale=3, model_pretrain=str(ck))).cuda().eval()
    x0 = torch.from_numpy(np.asarray(Image.open(tmp_path / "data" / "input" / "S01" / "low_light_10" / "00001.png")
                                     .convert("RGB").resize((W, H))).astype(np.float32) / 255.0).permute(2, 0, 1)[None]
    m.is_new_seq = True
    H2, H3, _ = m(x0.cuda())
    for t_, name in ((H3, "denoise"), (H2, "enhance")):
        want = np.clip(np.transpose(t_[0].cpu().numpy(), (1, 2, 0)) * 255.0, 0, 255.0).astype("uint8")
        got = np.asarray(Image.open(out / ("00001_%s.png" % name)))
        assert np.array_equal(got, want), name
    for t in range(1, n):
        im = np.asarray(Image.open(out / ("%05d_enhance.png" % (t + 1))))
        assert im.shape == (H, W, 3) and im.std() > 1.0

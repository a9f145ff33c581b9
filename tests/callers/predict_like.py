"""A caller written the way the reference's predict.py drives the model (predict.py:10-17,49-51,76-104), used by the
drop-in tests on the GPU box where /root/reference does not exist.  It is NOT a copy of predict.py: only the
interaction with ``model.model`` is reproduced -

  * ``from model.model import Finetunemodel`` resolved through ``sys.path`` (predict.py:10),
  * ``torch.set_default_device(cuda)`` BEFORE the model is built (predict.py:15-17),
  * ``Finetunemodel(args)`` with an argparse namespace, ``.to(device)``, ``.eval()``, ``requires_grad=False``,
  * a ``DataLoader(batch_size=1, pin_memory=True)`` yielding ``(frame, name, path, last_path)``, the
    ``is_new_seq`` rule of utils.sequential_judgment (utils/utils.py:145-160),
  * ``enhance, output, illum = model(input)`` under ``torch.no_grad()``,
  * ``clip(x*255, 0, 255).astype(uint8)`` and two PNGs per frame in ``<save>/<scene>/<brightness>/``.

A ``model/`` directory of its own sits beside this script (tests/callers/model/) and raises on import, like the
reference tree's own ``model`` package would shadow a PYTHONPATH entry: the test passes only when the launcher
(`python -m zero_tig_b200.run`) really puts the drop-in in front of the script directory."""
import argparse
import glob
import os
import sys

import numpy as np
import torch
import torch.utils.data
from PIL import Image
from model.model import Finetunemodel

device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
torch.set_default_device(device)

ap = argparse.ArgumentParser("predict-like")
ap.add_argument("--lowlight_images_path", required=True)
ap.add_argument("--save", required=True)
ap.add_argument("--model_pretrain", required=True)
ap.add_argument("--of_scale", type=int, default=3)
ap.add_argument("--dataset", default="RLV")
ap.add_argument("--size", default="400x390", help="WxH every frame is resized to (the reference uses 1920x1080)")
args = ap.parse_args()


class Frames(torch.utils.data.Dataset):
    def __init__(self, root, size):
        with open(os.path.join(root, "test_list.txt")) as f:
            scenes = [s.strip() for s in f if s.strip()]
        self.files = []
        for s in scenes:
            self.files += sorted(glob.glob(os.path.join(root, "input", s, "low_light_10", "*.png")))
        self.size = size
        self.last = self.files[0]

    def __len__(self):
        return len(self.files)

    def __getitem__(self, i):
        im = Image.open(self.files[i]).convert("RGB").resize(self.size)
        x = torch.from_numpy(np.asarray(im).astype(np.float32) / 255.0).permute(2, 0, 1).cpu()
        last, self.last = self.last, self.files[i]
        return x, os.path.splitext(os.path.basename(self.files[i]))[0], self.files[i], last


def new_sequence(path, last):
    d0, n0 = os.path.split(path)
    d1, n1 = os.path.split(last)
    return d0 != d1 or int(os.path.splitext(n0)[0]) != int(os.path.splitext(n1)[0]) + 1


def to_u8(t):
    return np.clip(np.transpose(t[0].cpu().float().numpy(), (1, 2, 0)) * 255.0, 0, 255.0).astype("uint8")


def main():
    w, h = [int(v) for v in args.size.split("x")]
    model = Finetunemodel(args).to(device).eval()
    for p in model.parameters():
        p.requires_grad = False
    import model.model as mm   # noqa: F811  (report which module the import resolved to)
    print("MODEL_MODULE_FILE=" + os.path.abspath(sys.modules["model.model"].__file__))
    data = Frames(args.lowlight_images_path, (w, h))
    queue = torch.utils.data.DataLoader(data, batch_size=1, pin_memory=True, num_workers=0, shuffle=False)
    with torch.no_grad():
        for x, name, path, last in queue:
            model.is_new_seq = new_sequence(path[0], last[0])
            enhance, output, illum = model(x.to(device))
            parts = path[0].split(os.sep)
            out_dir = os.path.join(args.save, parts[-3], parts[-2])
            os.makedirs(out_dir, exist_ok=True)
            Image.fromarray(to_u8(output)).save(os.path.join(out_dir, name[0] + "_denoise.png"), "PNG")
            Image.fromarray(to_u8(enhance)).save(os.path.join(out_dir, name[0] + "_enhance.png"), "PNG")
    print("FRAMES_DONE=%d" % len(data))


if __name__ == "__main__":
    main()

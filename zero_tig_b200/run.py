"""Launcher that runs an UNMODIFIED reference CLI on the B200 path.

    python -m zero_tig_b200.run /path/to/Zero-TIG/predict.py --model_pretrain weights/BVI-RLV.pt ...
    python -m zero_tig_b200.run /path/to/Zero-TIG/train.py   --lowlight_images_path data ...

``python script.py`` puts the script's directory at ``sys.path[0]``, ahead of PYTHONPATH, so the
reference's own ``model/`` package would always win.  This launcher reproduces the interpreter's
set-up (``sys.argv``, ``__main__``, the script directory on the path so ``dataloader``/``utils``/``loss``
resolve to the reference's files) but places ``zero_tig_b200/dropin`` in front of it, so
``from model.model import Finetunemodel`` (predict.py:10) / ``from model.model import *`` (train.py:12)
bind to this repository's classes.  Everything else of the CLI - flags, logging, output tree - is the
reference's code, unchanged.
"""
import os
import runpy
import sys


def install(script_dir=None):
    """Put the drop-in ``model`` package (and, after it, ``script_dir``) at the front of ``sys.path`` and
    forget any ``model`` package imported before.  Returns the drop-in directory."""
    here = os.path.dirname(os.path.abspath(__file__))
    dropin = os.path.join(here, "dropin")
    root = os.path.dirname(here)
    for p in (dropin, script_dir, root):
        while p and p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, dropin)
    if script_dir:
        sys.path.insert(1, script_dir)
    sys.path.append(root)
    for name in [m for m in sys.modules if m == "model" or m.startswith("model.")]:
        del sys.modules[name]
    return dropin


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help"):
        sys.stderr.write("usage: python -m zero_tig_b200.run <reference script.py> [script args...]\n")
        return 2
    script = os.path.abspath(argv[0])
    if not os.path.isfile(script):
        sys.stderr.write("zero_tig_b200.run: no such script: %s\n" % script)
        return 2
    install(os.path.dirname(script))
    sys.argv = [script] + argv[1:]
    runpy.run_path(script, run_name="__main__")
    return 0


if __name__ == "__main__":
    sys.exit(main())

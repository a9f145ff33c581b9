"""Directory that shadows the reference's top-level ``model`` package.

``zero_tig_b200/dropin/`` is meant to sit at ``sys.path[0]`` of an unmodified reference CLI run
(``predict.py:10`` ``from model.model import Finetunemodel``, ``train.py:12`` ``from model.model import *``);
``python -m zero_tig_b200.run <script> ...`` puts it there.  The ``model`` package inside locates the
repository root itself, so it also works when only ``PYTHONPATH=.../zero_tig_b200/dropin`` is set and the
script directory does not hold a ``model/`` of its own.
"""
import os

PATH = os.path.dirname(os.path.abspath(__file__))

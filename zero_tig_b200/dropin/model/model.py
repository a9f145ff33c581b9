"""``model.model`` as the reference CLIs import it (predict.py:10, train.py:12, evals.py).

``train.py`` does ``from model.model import *`` and then uses ``nn`` (``nn.utils.clip_grad_norm_``,
train.py:130) and ``np`` (train.py:58,69,134) from that star import, so the same module-level names the
reference's model/model.py:1-11 binds are bound here."""
import os.path  # noqa: F401

import numpy as np  # noqa: F401
import torch  # noqa: F401
import torch.nn as nn  # noqa: F401
import torch.nn.functional as F  # noqa: F401

from zero_tig_b200.model.model import (Denoise_1, Denoise_2, Enhancer, Finetunemodel,  # noqa: F401
                                       Network)
from zero_tig_b200.model.RAFT.raft import RAFT  # noqa: F401

"""RAFT parameter container + B200 forward.

Mirrors the module tree (and therefore the ``state_dict`` key set) of the reference
``model/RAFT/{raft,extractor,update}.py`` so shipped checkpoints load unchanged; the
forward pass does NOT run these torch modules - it hands their tensors to the
hand-written sm_100a kernels (``zero_tig_b200.engine.RaftEngine``).

Reference: RAFT.forward model/RAFT/raft.py:77-130; BasicEncoder extractor.py:117-191;
BasicUpdateBlock update.py:114-136.
"""
import math

import torch
import torch.nn as nn


def _norm(kind, ch):
    return nn.InstanceNorm2d(ch) if kind == "instance" else nn.BatchNorm2d(ch)


class ResidualBlock(nn.Module):
    """extractor.py:5-44 (only the 'instance' / 'batch' variants are reachable)."""

    def __init__(self, cin, cout, norm_fn, stride=1):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1, stride=stride)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        self.norm1 = _norm(norm_fn, cout)
        self.norm2 = _norm(norm_fn, cout)
        self.stride = stride
        self.downsample = None
        if stride != 1:
            self.norm3 = _norm(norm_fn, cout)
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride=stride), self.norm3)


class BasicEncoder(nn.Module):
    """extractor.py:117-156: 7x7/2 stem, 3 stages of 2 residual blocks, 1x1 head."""

    def __init__(self, output_dim=128, norm_fn="batch", dropout=0.0):
        super().__init__()
        self.norm_fn = norm_fn
        self.norm1 = _norm(norm_fn, 64)
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3)
        widths = [(64, 64, 1), (64, 96, 2), (96, 128, 2)]
        for i, (cin, cout, s) in enumerate(widths, start=1):
            setattr(self, f"layer{i}", nn.Sequential(ResidualBlock(cin, cout, norm_fn, s),
                                                     ResidualBlock(cout, cout, norm_fn, 1)))
        self.conv2 = nn.Conv2d(128, output_dim, 1)
        for m in self.modules():  # same init law as extractor.py:149-156
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)


class FlowHead(nn.Module):
    def __init__(self, input_dim=128, hidden_dim=256):
        super().__init__()
        self.conv1 = nn.Conv2d(input_dim, hidden_dim, 3, padding=1)
        self.conv2 = nn.Conv2d(hidden_dim, 2, 3, padding=1)


class SepConvGRU(nn.Module):
    def __init__(self, hidden_dim=128, input_dim=320):
        super().__init__()
        c = hidden_dim + input_dim
        for g in "zrq":
            setattr(self, f"conv{g}1", nn.Conv2d(c, hidden_dim, (1, 5), padding=(0, 2)))
        for g in "zrq":
            setattr(self, f"conv{g}2", nn.Conv2d(c, hidden_dim, (5, 1), padding=(2, 0)))


class BasicMotionEncoder(nn.Module):
    def __init__(self, corr_levels=4, corr_radius=4):
        super().__init__()
        planes = corr_levels * (2 * corr_radius + 1) ** 2
        self.convc1 = nn.Conv2d(planes, 256, 1)
        self.convc2 = nn.Conv2d(256, 192, 3, padding=1)
        self.convf1 = nn.Conv2d(2, 128, 7, padding=3)
        self.convf2 = nn.Conv2d(128, 64, 3, padding=1)
        self.conv = nn.Conv2d(64 + 192, 128 - 2, 3, padding=1)


class BasicUpdateBlock(nn.Module):
    def __init__(self, hidden_dim=128):
        super().__init__()
        self.encoder = BasicMotionEncoder()
        self.gru = SepConvGRU(hidden_dim, 128 + hidden_dim)
        self.flow_head = FlowHead(hidden_dim, 256)
        self.mask = nn.Sequential(nn.Conv2d(128, 256, 3, padding=1), nn.ReLU(inplace=True),
                                  nn.Conv2d(256, 64 * 9, 1))


class RAFT(nn.Module):
    """Drop-in for reference ``RAFT(args)`` (raft.py:23-48).  ``args`` is an
    ``argparse.Namespace``; the same defaults are written back into it."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.hidden_dim = self.context_dim = 128
        args.corr_levels, args.corr_radius = 4, 4
        for k, v in (("of_scale", 3), ("dropout", 0), ("alternate_corr", False), ("mixed_precision", False)):
            if k not in args:
                setattr(args, k, v)
        self.fnet = BasicEncoder(256, "instance")
        self.cnet = BasicEncoder(256, "batch")
        self.update_block = BasicUpdateBlock(128)
        self._engine = None

    def engine(self):
        from zero_tig_b200.engine import RaftEngine
        if self._engine is None:
            self._engine = RaftEngine(self)
        return self._engine

    def forward(self, image1, image2, iters=12, test_mode=True):
        """Returns (flow_low, flow_up) like raft.py:130 (flow_up at the padded size)."""
        return self.engine().forward(image1, image2, iters)

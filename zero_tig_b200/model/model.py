"""Drop-in replacements for the reference ``model/model.py`` classes.

Same constructor arguments, attributes, ``state_dict`` keys and return tuples as the
reference (``Finetunemodel`` model/model.py:262-384, ``Network`` :84-259), so ``predict.py``
/ ``train.py`` run unchanged with this directory first on ``sys.path``.  The modules only
hold parameters: all arithmetic runs in the sm_100a kernels of ``libzerotig.so`` through
``zero_tig_b200.engine`` - there is no torch-op or CPU fallback for the compute.
"""
import torch
import torch.nn as nn

from .RAFT.raft import RAFT


class Denoise_1(nn.Module):
    """3 -> 48 -> 48 -> 3 (3x3, 3x3, 1x1; LeakyReLU 0.2).  model/model.py:15-28."""

    def __init__(self, chan_embed=48, cin=3, cout=3):
        super().__init__()
        self.act = nn.LeakyReLU(negative_slope=0.2, inplace=True)
        self.conv1 = nn.Conv2d(cin, chan_embed, 3, padding=1)
        self.conv2 = nn.Conv2d(chan_embed, chan_embed, 3, padding=1)
        self.conv3 = nn.Conv2d(chan_embed, cout, 1)


class Denoise_2(Denoise_1):
    """12 -> 48 -> 48 -> 6.  model/model.py:31-44."""

    def __init__(self, chan_embed=96):
        super().__init__(chan_embed, 12, 6)


class Enhancer(nn.Module):
    """9 -> 64, three applications of ONE shared conv+BN+ReLU residual block, 64 -> 3 sigmoid.
    model/model.py:47-81 (``blocks`` aliases ``conv`` three times, so the state_dict carries
    ``enhance.conv.*`` and ``enhance.blocks.{0,1,2}.*`` for the same tensors)."""

    def __init__(self, layers, channels):
        super().__init__()
        self.in_conv = nn.Sequential(nn.Conv2d(9, channels, 3, 1, 1), nn.ReLU())
        self.conv = nn.Sequential(nn.Conv2d(channels, channels, 3, 1, 1), nn.BatchNorm2d(channels), nn.ReLU())
        self.blocks = nn.ModuleList([self.conv for _ in range(layers)])
        self.out_conv = nn.Sequential(nn.Conv2d(channels, 3, 3, 1, 1), nn.Sigmoid())


def _make_raft(args):
    raft = RAFT(args)
    raft.eval()
    for p in raft.parameters():
        p.requires_grad = False
    return raft


class _Recurrent:
    """State shared by both model classes (model.py:98-102,217-219)."""

    def _init_state(self):
        self.last_H3 = None
        self.last_H3_wp = None
        self.last_s3 = None
        self.last_s3_wp = None
        self.is_new_seq = True
        self._frame_engine = None

    def frame_engine(self):
        from zero_tig_b200.engine import FrameEngine
        if self._frame_engine is None:
            self._frame_engine = FrameEngine(self)
        return self._frame_engine

    def update_H3(self, H3, s3):
        self.last_H3 = H3.detach()
        self.last_s3 = s3.detach()

    def update_cache(self, last_H3, last_s3, L2):
        """RAFT flow last->this on the 1/of_scale frames, then warp both state planes
        (model.py:221-259 / :346-384).  Returns (warped_H3, warped_s3)."""
        with torch.no_grad():
            return self.frame_engine().update_cache(last_H3, last_s3, L2, self.of_scale)


class Finetunemodel(nn.Module, _Recurrent):
    """Inference model.  ``forward(input) -> (H2, H3, s3)``; model/model.py:262-344."""

    def __init__(self, args):
        super().__init__()
        self.enhance = Enhancer(layers=3, channels=64)
        self.denoise_1 = Denoise_1(chan_embed=48)
        self.denoise_2 = Denoise_2(chan_embed=48)
        # key-filtered merge of the checkpoint, exactly like model.py:271-277 (raft.* keys are
        # dropped because self.raft does not exist yet - SURVEY F4)
        loc = "cuda:0" if torch.cuda.is_available() else "cpu"
        pretrained = torch.load(args.model_pretrain, map_location=loc)
        own = self.state_dict()
        own.update({k: v for k, v in pretrained.items() if k in own})
        self.load_state_dict(own)
        self._init_state()
        self.raft = _make_raft(args)
        self.of_scale = args.of_scale

    def load_raft(self, args):
        return _make_raft(args)

    def forward(self, input):
        if self.enhance.training:
            # the reference's Finetunemodel in train() mode would normalise with batch statistics and update the
            # running ones; the inference engine folds the running statistics into the block convolution
            # (predict.py:78 always calls eval()).  Refuse instead of silently diverging.
            raise RuntimeError("Finetunemodel.forward runs eval-mode BatchNorm only: call model.eval() first "
                               "(train-mode statistics are the zero-shot training path, Network._loss)")
        with torch.no_grad():
            r = self.frame_engine().infer(input, self.is_new_seq, self.last_H3, self.last_s3, self.of_scale)
        self.last_H3_wp, self.last_s3_wp = r["H3_wp"], r["s3_wp"]
        self.update_H3(r["H3"], r["s3"])
        return r["H2"], r["H3"], r["s3"]

    def forward_graphed(self, input):
        """Beyond the reference interface (used by zero_tig_b200.pipeline): ``forward`` for a steady-state frame
        (``is_new_seq`` False) whose input lives in a STATIC device buffer, replayed as one CUDA graph.  Same
        results as ``forward``; the returned tensors are static buffers that the next call overwrites."""
        if self.enhance.training:
            raise RuntimeError("Finetunemodel.forward_graphed runs eval-mode BatchNorm only: call model.eval() first")
        if self.is_new_seq or self.last_H3 is None:
            return self.forward(input)
        with torch.no_grad():
            r = self.frame_engine().infer_graphed(input, self.last_H3, self.last_s3, self.of_scale)
        self.last_H3_wp, self.last_s3_wp = r["H3_wp"], r["s3_wp"]
        self.update_H3(r["H3"], r["s3"])
        return r["H2"], r["H3"], r["s3"]


class Network(nn.Module, _Recurrent):
    """Zero-shot training model (model/model.py:84-259): parameter tree, recurrent state,
    ``update_cache``, the 23-tensor training forward, the 17 loss terms and the backward pass of the
    zero-shot training step run on the CUDA kernels."""

    def __init__(self, args):
        super().__init__()
        self.enhance = Enhancer(layers=3, channels=64)
        self.denoise_1 = Denoise_1(chan_embed=48)
        self.denoise_2 = Denoise_2(chan_embed=48)
        self.is_WB = args.dataset == "underwater"
        self._init_state()
        self.raft = _make_raft(args)
        self.of_scale = args.of_scale

    def load_raft(self, args):
        return _make_raft(args)

    @staticmethod
    def _init_normal(m):
        if isinstance(m, nn.Conv2d):
            m.weight.data.normal_(0.0, 0.02)
            if m.bias is not None:
                m.bias.data.zero_()
        if isinstance(m, nn.BatchNorm2d):
            m.weight.data.normal_(1.0, 0.02)

    def enhance_weights_init(self, m):
        self._init_normal(m)

    def denoise_weights_init(self, m):
        self._init_normal(m)

    def frame_engine(self):
        from zero_tig_b200.engine import TrainEngine
        if self._frame_engine is None:
            self._frame_engine = TrainEngine(self)
        return self._frame_engine

    def forward(self, input):
        """23 tensors in the order of model/model.py:203.  BatchNorm follows ``self.training`` like the
        reference (train.py:116 / :138); RAFT always runs in eval mode (model.py:240)."""
        with torch.no_grad():
            outs = self.frame_engine().forward_train(input, self.is_new_seq, self.last_H3, self.last_s3,
                                                     self.of_scale, self.enhance.training)
        self.last_H3_wp, self.last_s3_wp = self.frame_engine().last_wp
        return outs

    def trainable(self):
        """The 20 trainable tensors in the order of engine.TRAIN_PARAMS."""
        from zero_tig_b200.engine import TRAIN_PARAMS
        named = dict(self.named_parameters())
        return [named[k] for k in TRAIN_PARAMS]

    def _loss(self, input):
        """LossFunction.forward over Network.forward (model.py:205-215).  With autograd enabled the returned
        scalar carries a graph node whose backward runs the CUDA backward pass and fills ``.grad`` of the 20
        trainable tensors, so ``loss.backward()`` / ``clip_grad_norm_`` / ``optimizer.step()`` of train.py:126-131
        work unchanged."""
        params = self.trainable()
        if torch.is_grad_enabled() and any(p.requires_grad for p in params):
            return _ZeroTigLoss.apply(self, input, *params)
        return self._loss_value(input)

    def _loss_value(self, input):
        outs = self(input)
        with torch.no_grad():
            terms = self.frame_engine().loss_terms(input, outs, self.is_WB)
        self.last_loss_terms = terms
        self.update_H3(outs[13], outs[14])
        return terms.sum().float()


class _ZeroTigLoss(torch.autograd.Function):
    """Graph node of ``Network._loss``: forward = CUDA forward + 17 loss terms, backward = the hand-written
    backward pass (engine.TrainEngine.backward).  The parameters are passed only so autograd routes the
    gradients to them."""

    @staticmethod
    def forward(ctx, net, input, *params):
        ctx.net = net
        return net._loss_value(input)

    @staticmethod
    def backward(ctx, grad_out):
        with torch.no_grad():
            grads = ctx.net.frame_engine().backward()
            grads = [g * grad_out.to(g.dtype) for g in grads]
        return (None, None, *grads)

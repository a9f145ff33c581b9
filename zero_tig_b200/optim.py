"""Fused optimiser step of the zero-shot training loop (train.py:98,126-131).

The reference does ``loss.backward(); clip_grad_norm_(model.parameters(), 5); optimizer.step()`` with
``torch.optim.Adam(lr=1e-4, betas=(0.9, 0.999), weight_decay=3e-4)``.  Those calls keep working on the drop-in
``Network`` (its ``_loss`` returns an autograd scalar).  ``ClipAdam`` is the B200 path for the same step: the
92 620 trainable values live in ONE flat fp32 buffer (the parameters become views of it), the gradients in
another, and one call does [all-reduce over the data-parallel group ->] global-norm clip -> Adam in two kernel
launches (csrc/train_bwd.cu: sumsq_kernel, clip_adam_kernel).  Data-parallel training needs exactly
one collective per step: a sum all-reduce of the flat gradient buffer (370 KB) over NCCL/NVLink; the 1/world
scale is folded into the clip kernel.
"""
import torch

from . import ops


class ClipAdam:
    def __init__(self, params, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=3e-4, max_norm=5.0,
                 process_group=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("ClipAdam: no trainable parameters")
        dev = self.params[0].device
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.group = process_group
        self.sizes = [p.numel() for p in self.params]
        n = sum(self.sizes)
        self.flat = torch.empty(n, device=dev, dtype=torch.float32)
        self.grad = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self.sumsq = torch.zeros(1, device=dev, dtype=torch.float64)
        self.step_count = 0
        off = 0
        for p, k in zip(self.params, self.sizes):   # parameters become views of the flat buffer
            self.flat[off:off + k].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + k].view_as(p.data)
            off += k
        self._attach_grads()

    def _attach_grads(self):
        off = 0
        for p, k in zip(self.params, self.sizes):
            p.grad = self.grad[off:off + k].view_as(p.data)
            off += k

    def zero_grad(self):
        """Zero the flat gradient buffer and (re)attach the per-parameter views (autograd accumulates in place)."""
        self.grad.zero_()
        self._attach_grads()

    def gather_grads(self):
        """Make sure the flat buffer holds the current gradients (covers callers that replaced ``.grad``)."""
        off = 0
        for p, k in zip(self.params, self.sizes):
            view = self.grad[off:off + k]
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad.reshape(-1))
            off += k

    def world_size(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(self.group)
        return 1

    def reduce_grads(self):
        """Data-parallel step: ONE sum all-reduce of the flat gradient buffer; returns the 1/world scale that
        the clip kernel applies."""
        world = self.world_size()
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(self.grad, op=dist.ReduceOp.SUM, group=self.group)
        return 1.0 / world

    @torch.no_grad()
    def step(self):
        self.gather_grads()
        scale = self.reduce_grads()
        self.step_count += 1
        ops.clip_adam(self.flat, self.grad, self.exp_avg, self.exp_avg_sq, self.sumsq, self.step_count, self.lr,
                      self.betas, self.eps, self.weight_decay, self.max_norm, scale)
        # the kernel wrote the parameters through raw pointers (torch's _version did not move): tell the engines'
        # packed-weight caches explicitly
        from .engine import bump_param_generation
        bump_param_generation(self.params)
        self._attach_grads()

    def grad_norm(self):
        """Global gradient norm seen by the last step (before clipping, after the 1/world scale)."""
        return float(torch.sqrt(self.sumsq)[0]) / self.world_size()

"""Host-buffer frame pipeline around ``Finetunemodel`` (the loop of predict.py:76-89 with the copies overlapped).

predict.py moves one frame to the GPU, runs the model, and copies ``H2`` / ``H3`` back for ``save_images``
(predict.py:57-61,86-104), all on one stream.  ``FramePipeline`` keeps that contract - host frames in, host
``(H2, H3)`` out, strictly in frame order, the recurrent state carried by the model - but runs the host->device
copy of frame t+1 and the device->host copy of frame t-1 on two side streams while frame t computes, and replays
every steady-state frame as ONE CUDA graph (``Finetunemodel.forward_graphed``).

Two byte-saving options move work that the reference does on the host onto the device without changing a bit of it:
``input_u8`` uploads the decoded 8-bit HWC image and applies ``transforms.ToTensor`` (value / 255,
multi_read_data.py:86-87,131) on the GPU; ``quantize_u8`` applies ``save_images`` (``clip(x*255, 0, 255)`` -> uint8,
predict.py:57-61) on the GPU and returns the two 8-bit HWC images.  Together a 1080p frame moves 18.7 MB over PCIe
instead of 74.6 MB.  No arithmetic of the model happens here: only copies, the two converters and stream events.
"""
import os

import torch

from . import ops


class FramePipeline:
    def __init__(self, model, H, W, depth=2, quantize_u8=False, input_u8=False, graph=True):
        if not torch.cuda.is_available():
            raise RuntimeError("FramePipeline needs a CUDA device (no CPU fallback)")
        self.model, self.H, self.W, self.depth = model, H, W, depth
        self.dev = next(model.parameters()).device
        self.quantize_u8, self.input_u8, self.graph = quantize_u8, input_u8, graph
        self.h2d, self.d2h = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)
        idt, ishape = (torch.uint8, (H, W, 3)) if input_u8 else (torch.float32, (1, 3, H, W))
        odt, oshape = (torch.uint8, (H, W, 3)) if quantize_u8 else (torch.float32, (1, 3, H, W))
        self.in_dtype, self.in_shape = idt, ishape
        self.dev_in = [torch.empty(ishape, dtype=idt, device=self.dev) for _ in range(depth)]
        self.x = torch.empty(1, 3, H, W, device=self.dev)                     # static model input (graph replay)
        self.dev_out = [[torch.empty(oshape, dtype=odt, device=self.dev) for _ in range(2)] for _ in range(depth)]
        self.host_out = [[torch.empty(oshape, dtype=odt, device="cpu").pin_memory() for _ in range(2)]
                         for _ in range(depth)]
        self.ev_in = [torch.cuda.Event() for _ in range(depth)]      # H2D of slot done
        self.ev_free = [torch.cuda.Event() for _ in range(depth)]    # compute finished reading the input slot
        self.ev_done = [torch.cuda.Event() for _ in range(depth)]    # compute finished writing the output slot
        self.ev_out = [torch.cuda.Event() for _ in range(depth)]     # D2H of slot done
        self.h2d_bytes = 3 * H * W * (1 if input_u8 else 4)
        self.d2h_bytes = 2 * 3 * H * W * (1 if quantize_u8 else 4)

    def _upload(self, slot, frame_host, first_use):
        if frame_host.dtype != self.in_dtype or tuple(frame_host.shape) != tuple(self.in_shape):
            raise ValueError("FramePipeline expects %s frames of shape %s, got %s %s" % (
                self.in_dtype, tuple(self.in_shape), frame_host.dtype, tuple(frame_host.shape)))
        with torch.cuda.stream(self.h2d):
            if not first_use:
                self.h2d.wait_event(self.ev_free[slot])
            self.dev_in[slot].copy_(frame_host, non_blocking=True)
            self.ev_in[slot].record(self.h2d)

    def run(self, frames, emit, new_sequence=True):
        """frames: sequence of pinned host tensors (one sequence, in order): fp32 [1,3,H,W], or uint8 [H,W,3] with
        ``input_u8``.  emit(i, H2_host, H3_host) is called in order once frame i's outputs are in host memory.  The two
        host tensors are pinned staging buffers that are valid ONLY for the duration of the callback: the device->host
        copy of a later frame is enqueued into the same buffers right after emit returns, so a consumer that keeps
        them (an asynchronous PNG writer, a list) must copy them first."""
        for _ in self.steps(frames, emit, new_sequence):
            pass

    def steps(self, frames, emit, new_sequence=True):
        """Generator form of ``run``: enqueues one frame per ``next()`` on the stream that is current at that
        moment (callers that interleave several pipelines keep one stream per pipeline, see ``MultiShardPipeline``)."""
        n = len(frames)
        if n == 0:
            return
        cur = torch.cuda.current_stream(self.dev)
        HW = self.H * self.W
        self._upload(0, frames[0], True)
        for i in range(n):
            slot = i % self.depth
            if i + 1 < n:
                self._upload((i + 1) % self.depth, frames[i + 1], i + 1 < self.depth)
            cur.wait_event(self.ev_in[slot])
            if self.input_u8:
                ops.u8_hwc_to_f32(self.dev_in[slot], self.x, 3, HW)      # ToTensor on the device
            else:
                self.x.copy_(self.dev_in[slot])
            self.ev_free[slot].record(cur)
            self.model.is_new_seq = bool(new_sequence and i == 0)
            if self.graph:
                H2, H3, _ = self.model.forward_graphed(self.x)
            else:
                H2, H3, _ = self.model(self.x)
            if i >= self.depth:                      # output slot about to be overwritten: hand it out first
                self.ev_out[slot].synchronize()
                emit(i - self.depth, *self.host_out[slot])
            if self.quantize_u8:                     # save_images on the device
                ops.quantize_u8_hwc(H2, self.dev_out[slot][0], 3, HW)
                ops.quantize_u8_hwc(H3, self.dev_out[slot][1], 3, HW)
            else:
                self.dev_out[slot][0].copy_(H2)
                self.dev_out[slot][1].copy_(H3)
            self.ev_done[slot].record(cur)
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(self.ev_done[slot])
                self.host_out[slot][0].copy_(self.dev_out[slot][0], non_blocking=True)
                self.host_out[slot][1].copy_(self.dev_out[slot][1], non_blocking=True)
                self.ev_out[slot].record(self.d2h)
            yield i
        for i in range(max(0, n - self.depth), n):
            slot = i % self.depth
            self.ev_out[slot].synchronize()
            emit(i, *self.host_out[slot])


def share_sms(models):
    """Throughput mode for S model replicas on one GPU: the persistent RAFT update kernel (csrc/raft_update.cu) is a
    cooperative launch that would otherwise claim all 148 SMs, so the chains of different shards could not overlap
    inside it; each replica's kernel is limited to 148 // S CTAs (measured at 1080p: 3 shards x 49 CTAs 248 frames/s,
    3 x 148 222 frames/s; a lone sequence runs fastest on the full grid).  A single replica keeps the full grid."""
    n = len(models)
    per = int(os.environ.get("ZT_SHARD_CTAS", "0")) or max(30, 148 // n)      # ZT_SHARD_CTAS: override for sweeps
    for m in models:
        eng = m.raft.engine()
        eng.FUSED_MAX_CTAS = 0 if n <= 1 else per


class MultiShardPipeline:
    """Several frame-range shards of a clip on ONE GPU (the same unit ``shard.plan`` deals to the GPUs of a box: a
    contiguous frame range preceded by one halo frame), each with its own model replica, recurrent state,
    ``FramePipeline`` and compute stream; frames are enqueued round-robin from one host thread.

    Why: a frame is ~250 dependent launches, most of them RAFT kernels on a 45x80 grid that cannot fill 148 SMs;
    a second, independent frame chain on another stream fills those holes.  Every shard computes exactly what a
    single ``FramePipeline`` over its (halo + range) sub-clip computes; what the cut does to the frames after it,
    compared with the unsharded run, is measured in tests/test_gpu_recurrence.py.
    """

    def __init__(self, models, H, W, depth=2, quantize_u8=False, input_u8=False, graph=True):
        if not models:
            raise ValueError("MultiShardPipeline needs at least one model replica")
        self.pipes = [FramePipeline(m, H, W, depth=depth, quantize_u8=quantize_u8, input_u8=input_u8, graph=graph)
                      for m in models]
        share_sms(models)
        self.dev = self.pipes[0].dev
        self.streams = [torch.cuda.Stream(self.dev) for _ in models]
        self.h2d_bytes, self.d2h_bytes = self.pipes[0].h2d_bytes, self.pipes[0].d2h_bytes

    def run(self, shard_frames, emit, new_sequence=True):
        """shard_frames[s]: the pinned host frames of shard s, in order (its halo frame first when ``new_sequence``).
        emit(s, i, H2_host, H3_host) is called in frame order within each shard (buffers valid inside the callback
        only, see ``FramePipeline.run``)."""
        if len(shard_frames) != len(self.pipes):
            raise ValueError("expected %d frame lists, got %d" % (len(self.pipes), len(shard_frames)))
        cur = torch.cuda.current_stream(self.dev)
        start = torch.cuda.Event()
        start.record(cur)
        gens = []
        for s, (pipe, frames) in enumerate(zip(self.pipes, shard_frames)):
            self.streams[s].wait_event(start)
            gens.append(pipe.steps(frames, (lambda i, a, b, _s=s: emit(_s, i, a, b)), new_sequence))
        live = list(range(len(gens)))
        while live:
            for s in list(live):
                with torch.cuda.stream(self.streams[s]):
                    try:
                        next(gens[s])
                    except StopIteration:
                        live.remove(s)
        for st in self.streams:                       # the caller's stream continues after every shard
            done = torch.cuda.Event()
            done.record(st)
            cur.wait_event(done)

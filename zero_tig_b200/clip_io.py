"""Threaded host I/O around the frame pipeline: PNG decode + resize in, two PNGs per frame out.

With the GPU path at hundreds of frames per second the reference's per-frame host work - ``Image.open(..).convert('RGB')
.resize((1920, 1080))`` + ``ToTensor`` in the dataloader (dataloader/multi_read_data.py:127-132) and two
``Image.fromarray(..).save(.., 'PNG')`` calls after the model (predict.py:57-61,101-104) - is what bounds the CLI.  This
module keeps those PIL calls UNCHANGED (same decoder, same default resize filter, same encoder settings, so the files are
byte-identical to what predict.py writes for the same model outputs) and moves them off the GPU's critical path:

* a pool of decode threads runs ahead of the GPU by a bounded number of frames and fills pinned 8-bit HWC buffers
  (``FramePipeline(input_u8=True)`` applies ToTensor's value / 255 on the device, bit-identical);
* the GPU quantises its outputs with ``save_images``' arithmetic on the device (``quantize_u8=True``, bit-identical) and
  a pool of encode threads writes the PNGs while later frames compute;
* sequences are cut exactly where predict.py cuts them (``sequential_judgment``, utils/utils.py:145-160).

PIL releases the GIL inside its decoder, resampler and zlib encoder, so plain threads scale across the host cores.
No arithmetic of the model happens here.
"""
import os
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

FRAME_SIZE = (1920, 1080)          # multi_read_data.py:129 (PIL order: width, height)


def is_new_sequence(img_path, last_img_path):
    """utils/utils.py:145-160: a frame starts a new sequence when its directory differs from the previous frame's or
    its integer file stem is not the previous stem + 1.  The dataloader hands the FIRST frame its own path as
    ``last_img_path`` (multi_read_data.py:96,138-139), which this rule also classifies as a new sequence."""
    img_dir, img_name = os.path.split(img_path)
    last_dir, last_name = os.path.split(last_img_path)
    if img_dir != last_dir:
        return True
    return int(os.path.splitext(img_name)[0]) != int(os.path.splitext(last_name)[0]) + 1


def split_sequences(paths):
    """[[paths of sequence 0], [paths of sequence 1], ...] in the order predict.py would process them."""
    seqs, last = [], None
    for p in paths:
        if last is None or is_new_sequence(p, last):
            seqs.append([])
        seqs[-1].append(p)
        last = p
    return seqs


def decode_frame(path, size=FRAME_SIZE):
    """RLVDataLoader.load_images_transform (multi_read_data.py:127-131) up to, not including, ToTensor:
    uint8 [H][W][3].  ``size`` = None keeps the file's own size (the other datasets' loaders do not resize)."""
    from PIL import Image
    im = Image.open(path).convert("RGB")
    if size is not None:
        im = im.resize(size)
    return np.array(im, dtype=np.uint8)                  # a writable copy (torch.from_numpy wants one)


def output_paths(img_path, save_root, dataset="RLV"):
    """(denoise_png, enhance_png) exactly as predict.py:90-104 names them."""
    name = os.path.splitext(os.path.basename(img_path))[0]
    if dataset == "RLV":
        parts = os.path.normpath(img_path).split(os.sep)
        save_dir = os.path.join(save_root, parts[-3], parts[-2])
    else:
        save_dir = os.path.join(save_root, os.path.basename(os.path.split(img_path)[0]))
    return os.path.join(save_dir, name + "_denoise.png"), os.path.join(save_dir, name + "_enhance.png")


def encode_png(arr, path):
    """predict.py:103-104: ``Image.fromarray(im).save(path, 'PNG')`` (default compression), directory created first."""
    from PIL import Image
    os.makedirs(os.path.dirname(path), exist_ok=True)
    Image.fromarray(arr).save(path, "PNG")


class _DecodedFrames:
    """List-like view of one sequence whose items are decoded by a thread pool at most ``ahead`` frames before they are
    asked for.  ``FramePipeline.steps`` indexes frames strictly in order (i, then i + 1), which is all this supports.

    With ``ring`` (a list of pre-allocated pinned uint8 [H][W][3] tensors, len >= ahead + 5) the decoders write into
    recycled pinned buffers - no allocation, and no CUDA call at all, on the worker threads (a ``pin_memory()`` there
    would invalidate a CUDA-graph capture running on the main thread).  Buffer n % len(ring) is reused for frame n
    only after frame n - len(ring) was emitted, i.e. after its host->device copy has completed (``released``)."""

    def __init__(self, paths, pool, size, ahead, ring=None):
        self.paths, self.pool, self.size, self.ahead, self.ring = paths, pool, size, max(1, ahead), ring
        self.futs, self.next_submit, self.released = {}, 0, -1
        if ring is not None and len(ring) < self.ahead + 5:
            raise ValueError("pinned ring too small for the look-ahead")

    def __len__(self):
        return len(self.paths)

    def release_upto(self, i):
        """Frames <= i have left the pipeline (called from emit)."""
        self.released = max(self.released, i)

    def _load(self, n):
        a = decode_frame(self.paths[n], self.size)
        if self.ring is None:
            return torch.from_numpy(a)
        dst = self.ring[n % len(self.ring)]
        if tuple(dst.shape) != a.shape:
            raise ValueError("frame %s is %s, the pipeline was built for %s" % (self.paths[n], a.shape, tuple(dst.shape)))
        dst.numpy()[...] = a
        return dst

    def _submit_until(self, hi):
        while self.next_submit < min(hi, len(self.paths)):
            n = self.next_submit
            if self.ring is not None and n - len(self.ring) > self.released:
                break                                 # its buffer is still owned by a frame in flight (cannot happen with
            self.futs[n] = self.pool.submit(self._load, n)     # len(ring) >= ahead + 5; kept as a guard)
            self.next_submit += 1

    def __getitem__(self, i):
        if i < 0 or i >= len(self.paths):
            raise IndexError(i)
        self._submit_until(i + 1 + self.ahead)
        if i not in self.futs:
            raise RuntimeError("frame %d requested before its pinned buffer was released" % i)
        for k in [k for k in self.futs if k < i - 1]:      # frames behind the pipeline's two-frame window are done with
            del self.futs[k]
        return self.futs[i].result()


class ClipRunner:
    """predict.py's loop (predict.py:75-104) with decode, GPU work and encode overlapped.

    ``process(frames, emit, new_sequence)`` is the engine: ``FramePipeline(model, H, W, quantize_u8=True,
    input_u8=True).run`` on the GPU box (``for_model`` builds it); any callable with that contract works, which is how
    the host logic is tested without a GPU.  ``emit`` receives frame-ordered ``(i, H2_u8, H3_u8)`` HWC uint8 tensors that
    are valid only inside the call (they are copied before the encoders see them)."""

    def __init__(self, process, save_root, dataset="RLV", size=FRAME_SIZE, decode_threads=None, encode_threads=None,
                 ahead=None, max_pending_writes=None):
        n = os.cpu_count() or 4
        self.process, self.save_root, self.dataset, self.size = process, save_root, dataset, size
        # the PNG encoder is ~20x the cost of decode + resize (tools/clip_io_bench.py): it gets the cores
        self.decode_threads = decode_threads or max(2, n // 4)
        self.encode_threads = encode_threads or max(2, n)
        self.ahead = ahead or 2 * self.decode_threads
        self.max_pending = max_pending_writes or 3 * self.encode_threads
        self.frame_shape = None                       # (H, W): set by for_model -> pinned ring buffers

    @classmethod
    def for_model(cls, model, save_root, H=FRAME_SIZE[1], W=FRAME_SIZE[0], dataset="RLV", resize=True, **kw):
        from .pipeline import FramePipeline
        pipe = FramePipeline(model, H, W, quantize_u8=True, input_u8=True)
        r = cls(pipe.run, save_root, dataset=dataset, size=(W, H) if resize else None, **kw)
        r.frame_shape = (H, W)
        return r

    def run(self, paths, skip=()):
        """Process ``paths`` (the dataloader's order) and write ``*_denoise.png`` / ``*_enhance.png`` for every frame.
        Frames whose path is in ``skip`` are computed (they advance the recurrent state) but not written: the halo
        frame of a frame-range shard (``shard.plan``).  Returns the number of frames written.  The first exception of
        any worker is re-raised here."""
        skip = frozenset(skip)
        written = [0]
        slots = threading.Semaphore(self.max_pending)
        lock = threading.Lock()
        errors = []

        def write_one(arr, path):
            try:
                encode_png(arr, path)
                with lock:
                    written[0] += 1
            except Exception as ex:                   # surfaced by run()
                errors.append(ex)
            finally:
                slots.release()

        ring = None
        if self.frame_shape is not None:              # pinned once, on this thread, before anything is captured
            ring = [torch.empty(self.frame_shape + (3,), dtype=torch.uint8).pin_memory() for _ in range(self.ahead + 5)]
        with ThreadPoolExecutor(self.decode_threads, thread_name_prefix="zt-decode") as dpool, \
                ThreadPoolExecutor(self.encode_threads, thread_name_prefix="zt-encode") as epool:
            for seq in split_sequences(list(paths)):
                frames = _DecodedFrames(seq, dpool, self.size, self.ahead, ring)

                def emit(i, h2, h3, _seq=seq, _frames=frames):
                    _frames.release_upto(i)
                    if errors:
                        raise errors[0]
                    if _seq[i] in skip:
                        return
                    d, e = output_paths(_seq[i], self.save_root, self.dataset)
                    for arr, path in ((h3, d), (h2, e)):     # predict.py: `output` (H3) -> *_denoise, `enhance` (H2) -> *_enhance
                        slots.acquire()               # bounds host memory when the encoders fall behind
                        epool.submit(write_one, arr.numpy().copy(), path)
                self.process(frames, emit, True)
        if errors:
            raise errors[0]
        return written[0] // 2


def rank_paths(paths, rank, world):
    """This rank's share of a dataloader-ordered frame list under ``shard.plan``: whole sequences first,
    contiguous frame ranges of a long sequence otherwise, each range that starts mid-sequence preceded by its halo
    frame.  Returns (paths to process, in order; halo paths to pass as ``skip``)."""
    from . import shard
    seqs = split_sequences(list(paths))
    todo, halos = [], []
    for w in shard.plan([len(s) for s in seqs], world)[rank]:
        if w.halo >= 0:
            todo.append(seqs[w.seq][w.halo])
            halos.append(seqs[w.seq][w.halo])
        todo.extend(seqs[w.seq][w.start:w.stop])
    return todo, halos


def predict_sharded(paths, save_root, make_runner, rank=None, world=None):
    """One rank of a frame-range-sharded prediction run (no collective: every rank reads its frames and writes its
    PNGs).  ``make_runner(save_root)`` builds this rank's ``ClipRunner`` (``ClipRunner.for_model`` on the GPU
    box); ``rank`` / ``world`` default to torchrun's RANK / WORLD_SIZE.  Note that two consecutive ranges of the same
    sequence stay two sequences for the runner: a range's halo is the frame before its first kept frame, so the path
    rule that cuts sequences sees consecutive indices inside a range and a fresh start at each range's first path."""
    rank = int(os.environ.get("RANK", "0")) if rank is None else rank
    world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else world
    todo, halos = rank_paths(paths, rank, world)
    return make_runner(save_root).run(todo, skip=halos)


"""Feature-shard dataset and pinned-memory loader for real training.

The split-inference pipelines dump the vision model's features with `torch.save` to `{feature_dir}/{seq}.h5`
(compressai_vision/pipelines/base.py:118-152, cfgs/generate_features.yaml:22-27): a dict
`{"data": {"p2": [N,256,H/4,W/4], "p3": ..., "p4": ..., "p5": ...}, "input_size": ..., "org_input_size": ...}`
(compressai_vision/model_wrappers/detectron2.py:96-102).  `FeatureFolder` reads such files and cuts aligned random
crops across the pyramid (the role `transforms.RandomCrop(patch_size)` plays for images in
compressai/examples/train.py:230-238); `ShardLoader` assembles batches as bf16 channels-last tensors in pinned host
memory on a background thread and copies them, on a copy stream, straight into the input buffers the captured train
step reads (`TrainStep.capture(n_buffers=2)`), so the H2D copy of batch i+1 overlaps step i -- the host side of what
`bench.py` measures as `e2e`.  Data-parallel ranks read disjoint strided subsets of the samples.
"""
import glob
import os
import queue
import threading
from typing import Dict, List, Tuple

import torch

from .spec import DEFAULT_SPEC, LMSFCSpec


def synth_features(spec: LMSFCSpec, B: int, H2: int, W2: int, seed: int = 1234, device="cpu", dtype=torch.float32):
    """The synthetic benchmark input: FPN-like features relu(N(0,1)) * s_l + N(0, 0.1), clipped to the
    dataset range [-26.43, 28.40] of compressai_vision/codecs/utils.py:30-39; seeded, identical on every rank count."""
    g = torch.Generator().manual_seed(seed)
    feats = {}
    for i, (k, (h, w)) in enumerate(zip(spec.levels, spec.level_shapes(H2, W2))):
        C = spec.channels[i]
        a = torch.relu(torch.randn(B, C, h, w, generator=g)) * (1.0 + 0.5 * i)
        a = a + 0.1 * torch.randn(B, C, h, w, generator=g)
        feats[k] = a.clamp_(-26.43, 28.40).to(device=device, dtype=dtype)
    return feats


class FeatureFolder:
    """Samples = (file, frame) pairs of the feature dumps under `root`.  `crop` is in INPUT pixels (256 -> a 64x64
    window of P2, 32x32 of P3, ...); it must be a multiple of 4 * 2^(levels-1) so the windows line up on every level."""

    def __init__(self, root: str, crop: Tuple[int, int] = (256, 256), spec: LMSFCSpec = DEFAULT_SPEC,
                 pattern: str = "*.h5", rank: int = 0, world: int = 1, seed: int = 1234, cache_files: int = 4):
        self.spec, self.crop = spec, (int(crop[0]), int(crop[1]))
        unit = spec.base_stride * 2 ** (len(spec.levels) - 1)
        if self.crop[0] % unit or self.crop[1] % unit:
            raise ValueError(f"crop {self.crop} must be a multiple of {unit} input pixels")
        self.files = sorted(glob.glob(os.path.join(root, pattern)))
        if not self.files:
            raise FileNotFoundError(f"no feature dumps matching {pattern} under {root}")
        self._cache: Dict[str, Dict[str, torch.Tensor]] = {}
        self._cache_order: List[str] = []
        self._cache_files = int(cache_files)
        samples = []
        for f in self.files:
            samples += [(f, i) for i in range(self._load(f)[spec.levels[0]].shape[0])]
        self.samples = samples[rank::world]          # disjoint per-rank shards (DistributedSampler's striding)
        if not self.samples:
            raise ValueError(f"rank {rank} of {world} has no samples ({len(samples)} in total)")
        self.gen = torch.Generator().manual_seed(seed + rank)

    def _load(self, path):
        if path not in self._cache:
            obj = torch.load(path, map_location="cpu")
            data = obj["data"] if isinstance(obj, dict) and "data" in obj else obj
            keys = list(data.keys())
            if len(keys) != len(self.spec.levels):
                raise ValueError(f"{path}: expected {len(self.spec.levels)} pyramid levels, got {keys}")
            lv = {}
            for name, k in zip(self.spec.levels, keys):      # levels are matched by ORDER, like pipelines/base.py:181-184
                t = data[k]
                t = torch.stack(list(t)) if isinstance(t, (list, tuple)) else t
                lv[name] = t if t.dim() == 4 else t.unsqueeze(0)
            self._cache[path] = lv
            self._cache_order.append(path)
            if len(self._cache_order) > self._cache_files:
                self._cache.pop(self._cache_order.pop(0), None)
        return self._cache[path]

    def __len__(self):
        return len(self.samples)

    def level_crop_shapes(self):
        h, w = self.crop[0] // self.spec.base_stride, self.crop[1] // self.spec.base_stride
        return self.spec.level_shapes(h, w)

    def crop_into(self, index: int, out: Dict[str, torch.Tensor], slot: int):
        """Writes one aligned random crop of sample `index` into out[level][slot] (any dtype / memory format)."""
        path, frame = self.samples[index % len(self.samples)]
        lv = self._load(path)
        top = lv[self.spec.levels[-1]]
        shapes = self.level_crop_shapes()
        th, tw = shapes[-1]
        # offsets are drawn on the top level's grid; ragged pyramids (P4 = 2*P5 - 1 rows, ...) bound them per level
        H = min(lv[name].shape[-2] >> d for d, name in enumerate(reversed(self.spec.levels)))
        W = min(lv[name].shape[-1] >> d for d, name in enumerate(reversed(self.spec.levels)))
        if H < th or W < tw:
            raise ValueError(f"{path}: pyramid {tuple(top.shape[-2:])} (top level) smaller than the crop's {th}x{tw}")
        y0 = int(torch.randint(0, H - th + 1, (1,), generator=self.gen))
        x0 = int(torch.randint(0, W - tw + 1, (1,), generator=self.gen))
        for d, (name, (h, w)) in enumerate(zip(reversed(self.spec.levels), reversed(shapes))):
            yy, xx = y0 << d, x0 << d
            src = lv[name][frame, :, yy:yy + h, xx:xx + w]
            if src.shape[-2:] != (h, w):
                raise ValueError(f"{path}: level {name} too small for an aligned {h}x{w} window at ({yy},{xx})")
            out[name][slot].copy_(src)
        return (y0, x0)


class ShardLoader:
    """Background-thread batcher into a ring of pinned bf16 channels-last staging batches."""

    def __init__(self, dataset: FeatureFolder, batch: int, ring: int = 4, pin: bool = None, shuffle: bool = True):
        self.ds, self.batch = dataset, int(batch)
        pin = torch.cuda.is_available() if pin is None else pin
        shapes = dataset.level_crop_shapes()
        chans = dataset.spec.channels

        def staging():
            b = {k: torch.empty((self.batch, c, h, w), dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
                 for k, c, (h, w) in zip(dataset.spec.levels, chans, shapes)}
            return {k: v.pin_memory() for k, v in b.items()} if pin else b
        self.ring = [staging() for _ in range(max(3, ring))]   # run_steps holds two while it waits for a third
        self.free: "queue.Queue[int]" = queue.Queue()
        self.full: "queue.Queue[int]" = queue.Queue()
        for i in range(len(self.ring)):
            self.free.put(i)
        self.shuffle = shuffle
        self._stop = False
        self.error = None
        self.thread = threading.Thread(target=self._work, daemon=True)
        self.thread.start()

    def _order(self):
        n = len(self.ds)
        return torch.randperm(n, generator=self.ds.gen).tolist() if self.shuffle else list(range(n))

    def _work(self):
        try:
            order, pos = self._order(), 0
            while not self._stop:
                try:
                    slot = self.free.get(timeout=0.1)
                except queue.Empty:
                    continue
                for s in range(self.batch):
                    if pos == len(order):
                        order, pos = self._order(), 0
                    self.ds.crop_into(order[pos], self.ring[slot], s)
                    pos += 1
                self.full.put(slot)
        except Exception as e:  # surfaced by next_host()
            self.error = e
            self.full.put(-1)

    def next_host(self):
        """-> (slot, {level: pinned bf16 NHWC batch}); give the slot back with release(slot) once it was consumed."""
        slot = self.full.get()
        if slot < 0:
            raise RuntimeError("ShardLoader worker failed") from self.error
        return slot, self.ring[slot]

    def release(self, slot: int):
        self.free.put(slot)

    def close(self):
        self._stop = True
        self.thread.join(timeout=5)

    def bytes_per_batch(self):
        return sum(v.numel() * v.element_size() for v in self.ring[0].values())


def run_steps(train_step, loader: ShardLoader, steps: int, log_every: int = 0):
    """Feeds a captured TrainStep (capture(..., n_buffers=2)) from `loader`: the H2D copy of batch i+1 runs on a copy
    stream into the buffer graph (i+1) % 2 reads while graph i % 2 runs.  Returns the [steps, 4] loss scalars
    (loss, bpp, mse, aux) read back through one pinned buffer."""
    dev = train_step.arena.device
    copy_stream = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream()
    bufs = [train_step.input_buffers(0), train_step.input_buffers(1)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    staged = [torch.cuda.Event(), torch.cuda.Event()]
    read = [torch.cuda.Event(), torch.cuda.Event()]
    read_stream = torch.cuda.Stream(device=dev)
    slots = [None, None]
    scalars = torch.zeros(steps, 4).pin_memory()
    for i in range(steps + 1):
        if i < steps:
            b = i % 2
            slot, host = loader.next_host()
            if slots[b] is not None:             # the staging batch copied two steps ago is free again
                staged[b].synchronize()
                loader.release(slots[b])
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[b])
                for k, v in host.items():
                    bufs[b][k].copy_(v, non_blocking=True)
                ready[b].record(copy_stream)
                staged[b].record(copy_stream)
            slots[b] = slot
        if i >= 1:
            b = (i - 1) % 2
            main.wait_event(ready[b])
            if i >= 3:
                main.wait_event(read[b])         # this graph's scalars of two steps ago have left the device
            out = train_step.replay(buffer=b)
            consumed[b].record(main)
            with torch.cuda.stream(read_stream):     # the 16-byte read-back on its own stream: the next replay does not
                read_stream.wait_event(consumed[b])  # queue behind a device-to-host copy
                scalars[i - 1].copy_(out["scalars"], non_blocking=True)
                read[b].record(read_stream)
            if log_every and i % log_every == 0:
                read[b].synchronize()
                print(f"step {i}: loss {scalars[i - 1, 0]:.4f} bpp {scalars[i - 1, 1]:.4f} mse {scalars[i - 1, 2]:.6f}")
    torch.cuda.synchronize()
    for b in (0, 1):
        if slots[b] is not None:
            loader.release(slots[b])
    return scalars

"""Data-parallel gradient exchange: one process per GPU, NCCL over NVLink 5 / NVSwitch.

Replaces nn.DataParallel of compressai/examples/train.py:64-71,275-276.  The gradient arena is
cut into contiguous buckets filled in reverse registration order (decoder first, the order
backward produces them); as soon as the last weight-gradient kernel of a bucket has been
*launched*, its all-reduce is enqueued on a side stream behind an event, so the exchange overlaps
the remaining dgrad / wgrad work.  `quantiles` (aux loss) are data independent and identical on
every rank: they are never reduced.
"""
import torch
import torch.distributed as dist


def owned_range(lo, hi, rank, world, align=64):
    """The part of arena slice [lo, hi) that `rank` owns (reduces, updates and publishes): `world` contiguous pieces,
    boundaries on multiples of `align` floats (256 bytes: float4 / TMA alignment); identical on every rank."""
    n = (hi - lo) // align                    # whole blocks; the tail block (if any) goes to the last rank
    a = lo + (n * rank // world) * align
    b = hi if rank == world - 1 else lo + (n * (rank + 1) // world) * align
    return a, b


class BucketedGradReducer:
    def __init__(self, model, bucket_mb=64.0, process_group=None, tail_mb=8.0):
        if model.arena is None:
            model.build_arena()
        self.model, self.arena = model, model.arena
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.grads = self.arena.net_grads
        self.comm_stream = torch.cuda.Stream(device=self.grads.device) if self.grads.is_cuda else None
        # ---- buckets: walk parameters from the END of the arena, close a bucket every bucket_mb
        cap = int(bucket_mb * (1 << 20) / 4)
        names = self.arena.names("net")
        self.buckets = []          # dicts {lo, hi, names(set), pending(int)}
        hi = self.grads.numel()
        cur_names, cur_lo = [], hi
        for n in reversed(names):
            cur_lo = self.arena.offsets[n]
            cur_names.append(n)
            if hi - cur_lo >= cap:
                self.buckets.append({"lo": cur_lo, "hi": hi, "names": set(cur_names)})
                hi, cur_names = cur_lo, []
        if cur_names:
            self.buckets.append({"lo": cur_lo, "hi": hi, "names": set(cur_names)})
        # The bucket at the START of the arena completes last (first conv layer, GDN group, EntropyBottleneck) and its
        # exchange cannot overlap anything: keep it small by cutting it at the parameter boundary that leaves at most
        # tail_mb in the final piece.
        tail_cap = int(tail_mb * (1 << 20) / 4)
        last = self.buckets[-1]
        if tail_cap > 0 and last["hi"] - last["lo"] > tail_cap:
            inside = sorted((self.arena.offsets[n], n) for n in last["names"])
            cut = None
            for o, n in inside:
                if o - last["lo"] <= tail_cap and o > last["lo"]:
                    cut = o
            if cut is not None:
                head = {n for o, n in inside if o >= cut}
                self.buckets[-1] = {"lo": cut, "hi": last["hi"], "names": head}
                self.buckets.append({"lo": last["lo"], "hi": cut, "names": last["names"] - head})
        self.bucket_of = {}
        for i, b in enumerate(self.buckets):
            for n in b["names"]:
                self.bucket_of[n] = i
        self.name_of = {id(p): n for n, p in self.arena.net}
        self._pending = None
        self._works = []
        self.launched = []
        # ---- hooks
        for m in model.modules():
            st = getattr(m, "_state", None)
            if st is not None and hasattr(st, "grad_ready"):
                st.grad_ready = self._on_ready
        self.eb_names = [n for n in names if ".entropy_bottleneck." in n or n.startswith("entropy_bottleneck.")]
        self.sync_replicas()

    def sync_replicas(self, optimizers=None):
        """What DistributedDataParallel does at construction: every rank starts from rank 0's parameters (and, on
        resume, optimiser state), whatever seed it built its model with.  The quantisation noise must NOT be shared:
        nn.DataParallel (examples/train.py:275-276) draws independent U(-.5,.5) per sample, so the rank is folded into
        the Philox key -- identical weights, distinct noise streams."""
        if self.world == 1 or not dist.is_initialized():
            return
        rank = dist.get_rank(self.pg)
        bufs = [b for b in (getattr(self.arena, "net_params", None), getattr(self.arena, "aux_params", None)) if b is not None]
        for opt in (optimizers or {}).values():
            bufs += [opt.m, opt.v, opt.step_count]
        for b in bufs:
            dist.broadcast(b, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0, group=self.pg)
        if self.grads.is_cuda:
            self.arena.refresh_shadows(cast=True)
            rng = self.model.entropy_bottleneck._rng(self.grads.device)
            if not getattr(rng, "_rank_folded", False):
                seed = (rng.seed + rank * 0x9E3779B97F4A7C15) & 0x7FFFFFFFFFFFFFFF
                rng.state[0] = seed
                rng.seed = seed
                rng._rank_folded = True

    # called at the start of every step (before backward)
    def begin(self):
        self._pending = [set(b["names"]) for b in self.buckets]
        self._works, self.launched = [], []

    def _on_ready(self, params):
        if self._pending is None:
            return
        for p in params:
            n = self.name_of.get(id(p))
            if n is None:
                continue
            i = self.bucket_of[n]
            pend = self._pending[i]
            pend.discard(n)
            if not pend and i not in self.launched:
                self._launch(i)

    def mark_ready(self, names):
        """For parameters whose gradients arrive through plain autograd (EntropyBottleneck)."""
        for n in names:
            i = self.bucket_of.get(n)
            if i is None or self._pending is None:
                continue
            self._pending[i].discard(n)
            if not self._pending[i] and i not in self.launched:
                self._launch(i)

    def _launch(self, i):
        self.launched.append(i)
        if self.world == 1:
            return
        b = self.buckets[i]
        view = self.grads[b["lo"]:b["hi"]]
        if self.comm_stream is not None:
            # a bucket's gradients may come from any of the weight-gradient streams (round-robin) and from the
            # current one: the exchange waits for what has been enqueued on each of them so far
            from .functional import SideStreams
            producers = {torch.cuda.current_stream()} | set(SideStreams._wgrad_pool)
            evs = []
            for s in producers:
                ev = torch.cuda.Event()
                ev.record(s)
                evs.append(ev)
            with torch.cuda.stream(self.comm_stream):
                for ev in evs:
                    self.comm_stream.wait_event(ev)
                dist.all_reduce(view, op=dist.ReduceOp.AVG, group=self.pg)
        else:   # gloo (CPU tests of the host logic): no AVG op, no streams
            w = dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
            self._works.append((w, view))

    def finish(self):
        """Flush buckets that never filled (unused parameters) and join the comm stream."""
        if self._pending is not None:
            for i in range(len(self.buckets)):
                if i not in self.launched:
                    self._launch(i)
        for w, view in self._works:
            w.wait()
            view.div_(self.world)
        if self.comm_stream is not None and self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._pending = None

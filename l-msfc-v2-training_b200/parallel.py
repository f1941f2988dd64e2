"""Data-parallel gradient exchange: one process per GPU, NCCL over NVLink 5 / NVSwitch.

Replaces nn.DataParallel of compressai/examples/train.py:64-71,275-276.  The gradient arena is
cut into contiguous buckets filled in reverse registration order (decoder first, the order
backward produces them); as soon as the last weight-gradient kernel of a bucket has been
*launched*, its all-reduce is enqueued on a side stream behind an event, so the exchange overlaps
the remaining dgrad / wgrad work.  `quantiles` (aux loss) are data independent and identical on
every rank: they are never reduced.
"""
import os

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
    def __init__(self, model, bucket_mb=64.0, process_group=None, tail_mb=8.0, edges=None):
        """`edges` (optional, ascending arena offsets starting at 0): explicit bucket boundaries instead of size-based ones."""
        if model.arena is None:
            model.build_arena()
        self.model, self.arena = model, model.arena
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.grads = self.arena.net_grads
        self.comm_stream = torch.cuda.Stream(device=self.grads.device) if self.grads.is_cuda else None
        if edges is not None:
            self._init_from_edges(edges)
            return
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
        self._finish_init()

    def _init_from_edges(self, edges):
        names = self.arena.names("net")
        edges = list(edges) + [self.grads.numel()]
        self.buckets = []
        for lo, hi in zip(edges[:-1], edges[1:]):
            inside = {n for n in names if lo <= self.arena.offsets[n] < hi}
            self.buckets.append({"lo": lo, "hi": hi, "names": inside})
        self._finish_init()

    def _finish_init(self):
        names = self.arena.names("net")
        self.bucket_of = {}
        for i, b in enumerate(self.buckets):
            for n in b["names"]:
                self.bucket_of[n] = i
        self.name_of = {id(p): n for n, p in self.arena.net}
        self._pending = None
        self._works = []
        self.launched = []
        # ---- hooks
        for m in self.model.modules():
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

    def _producer_events(self):
        """A bucket's gradients may come from any of the weight-gradient streams (round-robin) and from the current one:
        the exchange waits for what has been enqueued on each of them so far."""
        from .functional import SideStreams
        evs = []
        for s in {torch.cuda.current_stream()} | set(SideStreams._wgrad_pool):
            ev = torch.cuda.Event()
            ev.record(s)
            evs.append(ev)
        return evs

    def _launch(self, i):
        self.launched.append(i)
        if self.world == 1:
            return
        b = self.buckets[i]
        view = self.grads[b["lo"]:b["hi"]]
        if self.comm_stream is not None:
            evs = self._producer_events()
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


class SlicedNorm(BucketedGradReducer):
    """Single GPU: clip_grad_norm_'s squared norm (examples/train.py:100-101) slice by slice on a side stream, each part as
    soon as the slice's gradients are complete, instead of one 112 MB pass behind the last weight gradient (23 us of
    the 1 ms step with nothing to overlap).  Same readiness tracking as the multi-GPU reducers; fixed summation order."""

    def __init__(self, model, edges):
        from . import _lib as L
        self.L = L
        super().__init__(model, edges=edges)
        # high priority: a slice's norm is a few microseconds of work that must not queue behind a 148-CTA weight-gradient
        # launch of the NEXT slice (it trickled out on the step's tail otherwise)
        if os.environ.get("LMSFC_NORM_PRIO", "1") == "1":
            self.comm_stream = torch.cuda.Stream(device=self.grads.device, priority=-1)
        self.parts = torch.zeros(len(self.buckets), dtype=torch.float32, device=self.grads.device)
        self.outputs = None       # (sumsq, pending_flag or None): lets the LAST slice's launch add the parts up itself
        self._closed = False

    def sync_replicas(self, optimizers=None):
        pass

    def begin(self):
        self._closed = False
        return super().begin()

    def _launch(self, i):
        self.launched.append(i)
        b = self.buckets[i]
        evs = self._producer_events()
        with torch.cuda.stream(self.comm_stream):
            for ev in evs:
                self.comm_stream.wait_event(ev)
            st = torch.cuda.current_stream().cuda_stream
            if self.outputs is not None and len(self.launched) == len(self.buckets) and b["hi"] > b["lo"]:
                sumsq, flag = self.outputs            # every other part is complete on this stream
                self.L.sumsq_part_final(self.grads[b["lo"]:].data_ptr(), b["hi"] - b["lo"], self.parts.data_ptr(), i,
                                        len(self.buckets), sumsq.data_ptr(), None if flag is None else flag.data_ptr(), st)
                self._closed = True
            else:
                self.L.sumsq_part(self.grads[b["lo"]:].data_ptr(), b["hi"] - b["lo"], self.parts[i:].data_ptr(), st)

    def finish(self, sumsq, pending_flag=None):
        if self._pending is not None:
            for i in range(len(self.buckets)):
                if i not in self.launched:
                    self._launch(i)
        same = self.outputs is not None and self.outputs[0] is sumsq and self.outputs[1] is pending_flag
        if not (self._closed and same):
            with torch.cuda.stream(self.comm_stream):
                self.L.sum_parts(self.parts.data_ptr(), len(self.buckets), sumsq.data_ptr(),
                                 None if pending_flag is None else pending_flag.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._pending = None


# ====================================================================== the exchange as this library's own kernels
def symmetric_alloc(process_group=None):
    """-> alloc(numel, dtype, device) handing out peer-mapped ("symmetric") tensors of torch.distributed, and the list
    that collects (tensor, handle) pairs.  Plumbing only: the kernels get raw peer pointers (csrc/exchange.cu)."""
    import torch.distributed._symmetric_memory as symm
    group = process_group if process_group is not None else dist.group.WORLD
    made = []
    try:                                       # needed by older torch releases, a deprecated no-op in newer ones
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            symm.enable_symm_mem_for_group(group.group_name)
    except Exception:
        pass

    def alloc(numel, dtype, device):
        t = symm.empty(int(numel), dtype=dtype, device=device)
        try:
            hdl = symm.rendezvous(t, group.group_name)
        except TypeError:
            hdl = symm.rendezvous(t, group=group)
        t.zero_()
        made.append((t, hdl))
        return t
    alloc.made = made
    return alloc


class ShardedExchange(BucketedGradReducer):
    """Data-parallel step over NVLink peer memory with this library's own kernels (csrc/exchange.cu; replaces
    nn.DataParallel + clip_grad_norm_ + Adam of compressai/examples/train.py:64-71, 100-103, 275-276):

      * units = the arena slices of the pipelined train step (forward order); every rank OWNS 1/N of each unit;
      * backward: as soon as a unit's gradients are complete locally, `xchg_reduce` (exchange stream) waits for the
        same on every peer, then averages the owned part over the ranks' arenas with P2P loads in fixed rank order and
        reduces its squared norm deterministically -- overlapping the rest of backward;
      * `publish()` sends the rank's squared norm to every peer; `update()` waits for all N, so every rank derives the
        SAME clip coefficient, runs Adam on its owned parts only (1/N of the optimiser's HBM traffic per GPU) and stores
        the new bf16 operands (fp32 for GDN / entropy-bottleneck / bias parameters) into every rank's arena;
      * the forward pass waits per unit (`gate events`) for all ranks' stores.

    Replicas are bit-identical by construction (every operand is computed once, by its owner).  The fp32 master weights
    and Adam moments of a unit part live on its owner only: `consolidate()` gathers them (checkpoints)."""

    R_SLOT, NORM_SLOT, PUSH_SLOT = 0, 8, 16

    def __init__(self, model, edges, alloc, process_group=None):
        import os
        from . import _lib as L
        self.L = L
        super().__init__(model, process_group=process_group, edges=edges)
        arena = self.arena
        dev = arena.device
        self.rank = dist.get_rank(process_group)
        if self.world > L.MAX_RANKS:
            raise RuntimeError(f"the peer-memory exchange serves up to {L.MAX_RANKS} ranks of one node, got {self.world}")
        handle = {t.data_ptr(): h for t, h in alloc.made}
        try:
            hg, hp, hs = (handle[t.data_ptr()] for t in (arena._grads_all, arena.net_params, arena.net_bf16))
        except KeyError:
            raise RuntimeError("ShardedExchange: the model's arena was not built from symmetric memory (build_arena(alloc=...))")
        self.flags = alloc(L.XCHG_SLOTS * L.MAX_RANKS, torch.int64, dev)
        self.norm_all = alloc(L.MAX_RANKS, torch.float32, dev)
        hf, hn = handle_of(alloc, self.flags), handle_of(alloc, self.norm_all)
        self.tick = torch.zeros(1, dtype=torch.int64, device=dev)
        self.norm_parts = torch.zeros(len(self.buckets), dtype=torch.float32, device=dev)
        p = L.XchgPeers()
        p.world, p.rank = self.world, self.rank

        def peer_ptrs(h, t):
            """Rank k's address of tensor `t`: the handle lists allocation bases, `t` may sit at an offset inside."""
            delta = t.data_ptr() - int(h.buffer_ptrs[self.rank])
            mc = int(getattr(h, "multicast_ptr", 0) or 0)
            return [int(h.buffer_ptrs[k]) + delta for k in range(self.world)], (mc + delta if mc else 0)

        (pg_, mg), (pp_, mp), (ps_, ms) = peer_ptrs(hg, arena._grads_all), peer_ptrs(hp, arena.net_params), peer_ptrs(hs, arena.net_bf16)
        pf_, pn_ = peer_ptrs(hf, self.flags)[0], peer_ptrs(hn, self.norm_all)[0]
        for k in range(self.world):
            p.grads[k], p.params[k], p.shadow[k], p.flags[k], p.norm_all[k] = pg_[k], pp_[k], ps_[k], pf_[k], pn_[k]
        p.tick = self.tick.data_ptr()
        # NVSwitch multicast (NVLS) when the allocator provides it: in-switch reduction, one-store all-gather
        # measured (profiles/r2_scaling.txt): 2 ranks 1.096 ms/step with peer loads vs 1.105 with multicast, 4 ranks
        # 1.098 vs 1.056, 8 ranks 1.153 vs 1.040 -- the switch's reduction pays from 4 ranks on
        want_mc = os.environ.get("LMSFC_XCHG_MULTICAST", "1" if self.world > 2 else "0") == "1"
        use_mc = want_mc and mg and mp and ms
        self.multicast = bool(use_mc)
        if use_mc:
            p.mc_grads, p.mc_params, p.mc_shadow = mg, mp, ms
        self.peers = p
        self._handles = (hg, hp, hs, hf, hn)
        first_w = min(arena.offsets[n] for n, q in arena.net if q.dim() == 4)
        self.fp32_end = first_w                      # [0, fp32_end): parameters the kernels read in fp32
        self.units = [(b["lo"], b["hi"]) for b in self.buckets]
        self.own = [owned_range(lo, hi, self.rank, self.world) for lo, hi in self.units]
        self.reduce_ctas = int(os.environ.get("LMSFC_XCHG_REDUCE_CTAS", "0"))
        self.push_ctas = int(os.environ.get("LMSFC_XCHG_PUSH_CTAS", "0"))
        if os.environ.get("LMSFC_VERBOSE") and self.rank == 0:
            print(f"[lmsfc exchange] world {self.world}, units {self.units}, fp32_end {self.fp32_end}, "
                  f"multicast {self.multicast}", flush=True)
        if os.environ.get("LMSFC_XCHG_PRIO", "1") == "1":     # exchange kernels ahead of queued weight-gradient CTAs
            self.comm_stream = torch.cuda.Stream(device=dev, priority=-1)
        L.xchg_init()
        dist.barrier(group=process_group)            # every rank's buffers are mapped and zeroed before the first signal

    # ---- backward side
    def _launch(self, i):
        self.launched.append(i)
        evs = self._producer_events()
        a, b = self.own[i]
        with torch.cuda.stream(self.comm_stream):
            for ev in evs:
                self.comm_stream.wait_event(ev)
            self.L.xchg_reduce(self.peers, self.R_SLOT + i, a, b, self.norm_parts[i:].data_ptr(), self.reduce_ctas,
                               torch.cuda.current_stream().cuda_stream)

    def finish(self, pending_flag=None):
        """Flush units that never reported, then publish this rank's squared gradient norm (raises `pending_flag`)."""
        if self._pending is not None:
            for i in range(len(self.buckets)):
                if i not in self.launched:
                    self._launch(i)
        with torch.cuda.stream(self.comm_stream):
            self.L.xchg_norm_publish(self.peers, self.norm_parts.data_ptr(), len(self.buckets), self.NORM_SLOT,
                                     None if pending_flag is None else pending_flag.data_ptr(),
                                     torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._pending = None

    # ---- update side (current stream)
    def norm_wait(self, sumsq):
        self.L.xchg_norm_wait(self.peers, self.NORM_SLOT, sumsq.data_ptr(), torch.cuda.current_stream().cuda_stream)

    def adam_push(self, i, opt):
        lo, hi = self.units[i]
        a, b = self.own[i]
        b1, b2 = opt.betas
        self.L.xchg_adam_push(self.peers, self.PUSH_SLOT + i, lo, hi, a, b, self.fp32_end, opt.m.data_ptr(), opt.v.data_ptr(),
                              b1, b2, opt.eps, opt.scratch.data_ptr(), self.push_ctas, torch.cuda.current_stream().cuda_stream)

    def wait_pushed(self, i):
        self.L.xchg_wait(self.peers, self.PUSH_SLOT + i, torch.cuda.current_stream().cuda_stream)

    def update(self, opt, after_first=None):
        """The whole update in the eager order: norm -> clip / Adam on the owned parts -> wait for everybody's stores."""
        self.norm_wait(opt.sumsq)
        opt.prep()
        for i in range(len(self.units)):
            self.adam_push(i, opt)
        for i in range(len(self.units)):
            self.wait_pushed(i)
            if i == 0 and after_first is not None:
                after_first()

    @torch.no_grad()
    def consolidate(self, optimizers=()):
        """Gather the owners' fp32 master weights (and Adam moments) so that every rank holds the full state."""
        for i, (lo, hi) in enumerate(self.units):
            for k in range(self.world):
                a, b = owned_range(lo, hi, k, self.world)
                if b <= a:
                    continue
                src = dist.get_global_rank(self.pg, k) if self.pg is not None else k
                for buf in [self.arena.net_params] + [t for o in optimizers for t in (o.m, o.v)]:
                    dist.broadcast(buf[a:b], src=src, group=self.pg)


def handle_of(alloc, t):
    for u, h in alloc.made:
        if u.data_ptr() == t.data_ptr():
            return h
    raise KeyError("not a symmetric tensor")

"""The L-MSFC train step: one iteration of train_one_epoch (compressai/examples/train.py:85-118)

    zero_grad x2 -> forward -> RD loss -> backward -> [DP all-reduce] -> clip(1.0) + Adam(1e-4)
    -> aux loss -> backward -> Adam(1e-3) on quantiles   (Adam also writes the bf16 GEMM operands)

`TrainStep.step(feats)` runs it eagerly; `TrainStep.capture(feats)` records the same launch sequence
once into a CUDA graph (static shapes, device-side RNG counter and Adam step counter) so later
steps are a single graph launch with no Python or launch latency on the critical path.

`capture(feats, pipelined=True)` records the SAME iteration rotated by half a step:

    [clip + Adam of the previous backward pass, slice by slice on its own stream  ||  forward] -> RD loss
    -> backward -> [DP all-reduce] -> squared gradient norm (kept on the device for the next replay)

Every replay still does exactly one forward, one backward and one optimiser update, on the same weights in
the same order as the eager loop (forward k sees the weights after k updates); only the graph boundary moves,
so the HBM-bound update hides under the tensor-bound analysis convolutions.  The forward pass waits per slice
(model.LMSFC._gate), in the order it consumes the weights.  `flush()` applies the update still pending after
the last replay (call it before reading / saving the weights).
"""
import torch

from . import _lib as L
from . import functional as Fn
from .losses import RateDistortionLoss
from .model import LMSFC
from .optim import net_aux_optimizer
from .parallel import BucketedGradReducer, ShardedExchange, SlicedNorm, symmetric_alloc


class TrainStep:
    def __init__(self, model: LMSFC, lmbda=0.0130, lr=1e-4, aux_lr=1e-3, clip_max_norm=1.0, distributed=False,
                 bucket_mb=64.0, side_streams=True, exchange=None):
        """distributed: one process per GPU (torch.distributed initialised).  exchange: "p2p" (default; env LMSFC_DP) --
        this library's own reduce-scatter / sharded Adam / all-gather kernels over peer-mapped memory
        (parallel.ShardedExchange, csrc/exchange.cu) -- or "nccl": bucketed all-reduce + full Adam on every rank."""
        import os
        self.model = model
        self.sharded = False
        alloc = None
        if distributed:
            exchange = exchange or os.environ.get("LMSFC_DP", "p2p")
            if exchange not in ("p2p", "nccl"):
                raise ValueError(f'exchange must be "p2p" or "nccl", got {exchange!r}')
            if exchange == "p2p":
                try:
                    alloc = symmetric_alloc()
                    model.build_arena(alloc=alloc)   # (re)home parameters, gradients and bf16 operands in peer-mapped memory
                    self.sharded = True
                except Exception as e:               # no peer-mappable memory (e.g. GPUs without P2P): NCCL all-reduce
                    import warnings
                    warnings.warn(f"peer-memory exchange unavailable ({e}); falling back to the NCCL all-reduce")
                    alloc = None
                    model.arena = None
        if model.arena is None:
            model.build_arena()
        self.arena = model.arena
        self.criterion = RateDistortionLoss(lmbda=lmbda, spec=model.spec)
        self.opt = net_aux_optimizer(model, {"net": {"type": "Adam", "lr": lr}, "aux": {"type": "Adam", "lr": aux_lr}},
                                     clip_max_norm=clip_max_norm)
        # pipelined order: "the gradients of the last backward pass still await their update", kept on the device
        # (raised by the gradient-norm kernel, consumed by the update's prep kernel) so that a captured replay after
        # flush() applies nothing; the aux optimiser follows the net optimiser's decision
        self.pending_flag = torch.zeros(1, dtype=torch.int32, device=self.arena.device)
        self._one = torch.ones((), dtype=torch.float32, device=self.arena.device)
        self.opt["net"].pending_flag = self.pending_flag
        self.opt["aux"].gate = self.opt["net"]
        if self.sharded:
            self.reducer = ShardedExchange(model, [lo for lo, _, _ in self._slices()], alloc)
        else:
            self.reducer = BucketedGradReducer(model, bucket_mb=bucket_mb) if distributed else None
        if self.reducer is not None:
            self.reducer.sync_replicas(self.opt)     # replicas start from rank 0's weights, with distinct noise streams
        # single GPU, pipelined order: the gradient norm slice by slice beside the backward pass
        self.norms = (SlicedNorm(model, [lo for lo, _, _ in self._slices()])
                      if (self.reducer is None and clip_max_norm > 0 and os.environ.get("LMSFC_SLICED_NORM", "1") == "1") else None)
        if self.norms is not None:       # the slice that completes last adds the parts up in its own launch
            self.norms.outputs = (self.opt["net"].sumsq, self.pending_flag)
        if side_streams:
            Fn.SideStreams.enable(self.arena.device)
        self.gate_stream = torch.cuda.Stream(device=self.arena.device, priority=-1) if self.sharded else None
        self.graph = None
        self.static_in = None
        self.static_out = None
        self.launches_per_step = None
        self.pipelined = False
        self.pending = False        # pipelined mode: gradients (and their norm) of the last backward await their update

    # ------------------------------------------------------------------ one eager iteration
    def _iteration(self, feats, noise_z=None, noise_y=None):
        model, opt = self.model, self.opt
        opt["net"].zero_grad()
        opt["aux"].zero_grad()
        if self.reducer is not None:
            self.reducer.begin()
        out = model(feats, noise_z, noise_y)
        crit = self.criterion(out, feats)
        crit["loss"].backward()
        if self.reducer is not None:
            self.reducer.mark_ready(self.reducer.eb_names)
            self.reducer.finish()
        if self.sharded:                        # norm exchange -> clip + Adam on the owned parts -> operands to every rank
            self.reducer.update(opt["net"], after_first=self.arena.prepare)
        else:
            opt["net"].step()                   # clip_grad_norm_ + Adam, bf16 shadow in the same pass
        aux = model.aux_loss()
        aux.backward()
        opt["aux"].step()
        return {"loss": crit["loss"].detach(), "mse_loss": crit["mse_loss"].detach(),
                "bpp_loss": crit["bpp_loss"].detach(), "aux_loss": aux.detach()}

    def step(self, feats, noise_z=None, noise_y=None):
        if self.graph is not None and noise_z is None and noise_y is None:
            return self.replay(feats)
        self.flush()
        return self._iteration(feats, noise_z, noise_y)

    # ------------------------------------------------------------------ the iteration rotated by half a step
    def _slices(self):
        """Arena slices in the order the forward pass consumes them: [GDN group, EntropyBottleneck, g_a.conv2],
        [g_a.conv3], [g_a.conv4], [g_a.conv5], [h_a, h_s], [g_s].  Each slice opens a gate the forward pass waits on
        (model.LMSFC.GATES)."""
        off, n = self.arena.offsets, self.arena.n_net
        names = ["start"] + [k for k in self.model.GATES if k in off]
        edges = [0] + [off[k] for k in names[1:]] + [n]
        assert edges == sorted(edges), "arena order no longer follows the forward pass"
        return [(edges[i], edges[i + 1], names[i]) for i in range(len(names))]

    def _apply_pending(self):
        """clip + Adam of the pending gradients (on SideStreams.opt when enabled), gradient cleared in the same pass;
        then the aux step (examples/train.py:104-107).  Returns the aux loss."""
        model, opt, arena = self.model, self.opt, self.arena
        net, aux_opt = opt["net"], opt["aux"]
        main = torch.cuda.current_stream()
        os_ = Fn.SideStreams.opt
        ctx = torch.cuda.stream(os_) if os_ is not None else _Null()
        if os_ is not None:
            os_.wait_stream(main)
        arena.gates = {}
        with ctx:
            if self.sharded:
                self.reducer.norm_wait(net.sumsq)
            net.prep(use_pending=True)
            gs = self.gate_stream if (self.sharded and os_ is not None) else None
            for i, (lo, hi, gate) in enumerate(self._slices()):
                if self.sharded:               # Adam on the owned part, its operands stored into every rank's arena
                    self.reducer.adam_push(i, net)
                else:
                    net.update(lo, hi, zero_grad=True)
                if gs is not None:             # "every rank's part of the slice has landed" is waited for on a stream of
                    pushed = torch.cuda.Event()   # its own: this rank's next slice does not wait for the slowest peer
                    pushed.record()
                    gs.wait_event(pushed)
                with (torch.cuda.stream(gs) if gs is not None else _Null()):
                    if self.sharded:
                        self.reducer.wait_pushed(i)
                    if i == 0:
                        if os_ is not None:                        # "start": the first convolution needs its weights only,
                            ev = torch.cuda.Event()                # not the reparametrised GDN operands
                            ev.record()
                            arena.gates[gate] = ev
                        arena._grads_all[arena.n_net:].zero_()     # GDN effective-gradient scratch
                        arena.prepare()                            # GDN group (first in the arena) -> effective beta / gamma
                        if os_ is not None:                        # waited for in front of the first GDN launch
                            ev = torch.cuda.Event()                # (functional.wait_gdn_gate)
                            ev.record()
                            arena.gates["gdn"] = ev
                    elif os_ is not None:
                        ev = torch.cuda.Event()
                        ev.record()
                        arena.gates[gate] = ev
            if gs is not None:
                os_.wait_stream(gs)
            aux_opt.zero_grad()
            aux = model.aux_loss()
            aux.backward()
            aux_opt.prep(use_pending=True)
            aux_opt.update(0, aux_opt.flat_params.numel())
        if os_ is not None:
            main.wait_event(arena.gates.pop("start"))          # the later gates are waited on inside model.forward
        self.pending = False
        return aux.detach()

    def _grad_pass(self, feats, aux=None):
        """forward -> RD loss -> backward -> [DP all-reduce] -> squared gradient norm; leaves the update pending.
        aux (the aux loss of this replay's update): the four loss scalars are packed right behind the loss kernels, not
        behind the backward pass (where the tiny copy kernel would sit on the step's critical tail)."""
        model, opt = self.model, self.opt
        if self.reducer is not None:
            self.reducer.begin()
        if self.norms is not None:
            self.norms.begin()
        out = model(feats)
        if Fn.SideStreams.opt is not None:       # gradient buffers are cleared by the update pass: join before backward
            torch.cuda.current_stream().wait_stream(Fn.SideStreams.opt)
        self.arena.gates = {}
        crit = self.criterion(out, feats, extra=aux)
        if aux is not None and "scalars" not in crit:
            crit["scalars"] = torch.stack([crit["loss"].detach(), crit["bpp_loss"].detach(), crit["mse_loss"].detach(), aux])
        crit["loss"].backward(self._one)          # a kept 1.0: no ones_like fill between the loss and the first dgrad
        if self.reducer is not None:
            self.reducer.mark_ready(self.reducer.eb_names)
            if self.sharded:                     # publishes this rank's squared norm and raises the pending flag
                self.reducer.finish(pending_flag=self.pending_flag)
            else:
                self.reducer.finish()
        if self.norms is not None:
            self.norms.mark_ready(self.norms.eb_names)
            self.norms.finish(opt["net"].sumsq, self.pending_flag)
        elif not self.sharded:
            opt["net"].accumulate_norm()
        self.pending = True
        return crit

    def _iteration_pipelined(self, feats):
        aux = self._apply_pending()
        crit = self._grad_pass(feats, aux)
        return {"loss": crit["loss"].detach(), "mse_loss": crit["mse_loss"].detach(),
                "bpp_loss": crit["bpp_loss"].detach(), "aux_loss": aux, "scalars": crit["scalars"]}

    def flush(self):
        """Apply the update left pending by the last pipelined replay."""
        if self.pending:
            self._apply_pending()
            if Fn.SideStreams.opt is not None:
                torch.cuda.current_stream().wait_stream(Fn.SideStreams.opt)
            self.arena.gates = {}

    # ------------------------------------------------------------------ CUDA graph
    def capture(self, feats, warmup=2, pipelined=False, n_buffers=1, resume=False):
        """Record one iteration.  `feats` fixes the shapes; its values are copied into static
        input buffers on every replay.  pipelined: see the module docstring (call flush() at the end).
        n_buffers > 1 records the same iteration once per input buffer (`input_buffers(k)`, `replay(buffer=k)`):
        a loader copies host features straight into buffer k+1 while the step on buffer k runs -- no staging copy.
        The eager warm-up iterations before the capture TRAIN (they are real steps); resume=True (after
        load_state_dict) restores weights, optimiser state and the noise counter afterwards, so the first replay
        continues the saved run exactly."""
        lv = self.model.spec.levels
        dev = self.arena.device
        snap = None
        if resume:
            self.flush()
            rng = self.model.entropy_bottleneck._rng(dev)
            keep = [self.arena.net_params, self.arena.aux_params, rng.state]
            for o in self.opt.values():
                keep += [o.m, o.v, o.step_count]
            snap = [(t, t.clone()) for t in keep]
            warmup = max(1, int(warmup))         # at least one eager pass: lazy one-time initialisation cannot be captured
        self.static_ins = [{k: Fn.nhwc(feats[k], Fn.BF16).clone(memory_format=torch.channels_last) for k in lv}
                           for _ in range(max(1, int(n_buffers)))]
        self.static_in = self.static_ins[0]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):             # allocator / lazy-init warm-up outside capture
                self._iteration(self.static_in)
            if pipelined:                       # prime the pipeline: gradients of one pass, update pending
                self._iteration_pipelined_warm(self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.pipelined = bool(pipelined)
        if snap is not None:                     # undo the warm-up's training; nothing is pending
            with torch.no_grad():
                for t, saved in snap:
                    t.copy_(saved)
            self.arena.zero_grad()
            self.pending_flag.zero_()
            self.pending = False
            self.arena.refresh_shadows(cast=True)
            torch.cuda.synchronize()
        # the critical path (forward / data-gradient chain) is captured on a high-priority stream: its CTAs are
        # dispatched before those of the weight-gradient / optimiser / hyper / head streams whenever SMs free up
        cap = torch.cuda.Stream(device=dev, priority=-1)
        self.graphs, self.static_outs = [], []
        for buf in self.static_ins:
            c0 = L.launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap):
                out = (self._iteration_pipelined if pipelined else self._iteration)(buf)
                # the four loss scalars packed for one device-to-host read per step
                if "scalars" not in out:
                    out["scalars"] = torch.stack([out["loss"], out["bpp_loss"], out["mse_loss"], out["aux_loss"]])
            self.launches_per_step = L.launch_count() - c0
            self.graphs.append(g)
            self.static_outs.append(out)
        self.graph, self.static_out = self.graphs[0], self.static_outs[0]
        if snap is not None:
            self.pending = False                 # the capture recorded, not ran, a backward pass
        return self

    def release(self):
        """Drop the captured graphs (before destroying an NCCL communicator whose kernels they hold)."""
        self.graphs, self.static_outs = [], []
        self.graph = self.static_out = None

    def input_buffers(self, k=0):
        """The static bf16 NHWC input tensors graph k reads ({level: tensor}); fill them before replay(buffer=k)."""
        return self.static_ins[k]

    def _iteration_pipelined_warm(self, feats):
        """Eager rotated iterations before capture: allocator warm-up of the rotated order, and the pipeline primed."""
        self.opt["net"].zero_grad()
        self._grad_pass(feats)
        self._iteration_pipelined(feats)

    def replay(self, feats=None, buffer=0):
        if feats is not None:
            for k, buf in self.static_ins[buffer].items():
                src = feats[k]
                if src.data_ptr() != buf.data_ptr():
                    buf.copy_(src, non_blocking=True)
        for o in self.opt.values():
            o.sync_lr()                          # an LR scheduler may have written param_groups[..]["lr"]
        self.graphs[buffer].replay()
        if self.pipelined:
            self.pending = True                  # the replay's backward pass awaits its update (flush())
        return self.static_outs[buffer]

    # ------------------------------------------------------------------ checkpoint / resume (examples/train.py:283-321)
    def state_dict(self):
        """Everything the step keeps besides the model's own state_dict: both optimisers (Adam moments, step counters,
        learning rates) and the device-side noise counter.  A pending pipelined update is applied first."""
        self.consolidate()
        rng = self.model.entropy_bottleneck._rng(self.arena.device)
        return {"optimizer": self.opt["net"].state_dict(), "aux_optimizer": self.opt["aux"].state_dict(),
                "rng_state": rng.state.clone(), "lmbda": self.criterion.lmbda}

    def consolidate(self):
        """Apply a pending update and, under the sharded exchange, gather the owners' fp32 master weights and Adam
        moments, so that this rank's model / optimiser state is complete (collective: call on every rank)."""
        self.flush()
        if self.sharded:
            self.reducer.consolidate([self.opt["net"]])
            torch.cuda.synchronize()

    def load_state_dict(self, sd):
        """Restore after `model.load_state_dict(...)`; the next step continues the saved run."""
        self.flush()
        self.opt["net"].load_state_dict(sd["optimizer"])
        self.opt["aux"].load_state_dict(sd["aux_optimizer"])
        rng = self.model.entropy_bottleneck._rng(self.arena.device)
        rng.state.copy_(sd["rng_state"])
        self.pending_flag.zero_()
        self.pending = False
        self.model.refresh_shadows()


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

"""The L-MSFC train step: one iteration of train_one_epoch (compressai/examples/train.py:85-118)

    zero_grad x2 -> forward -> RD loss -> backward -> [DP all-reduce] -> clip(1.0) + Adam(1e-4)
    -> aux loss -> backward -> Adam(1e-3) on quantiles   (Adam also writes the bf16 GEMM operands)

`TrainStep.step(feats)` runs it eagerly; `TrainStep.capture(feats)` records the same launch sequence
once into a CUDA graph (static shapes, device-side RNG counter and Adam step counter) so later
steps are a single graph launch with no Python or launch latency on the critical path.
"""
import torch

from . import functional as Fn
from .losses import RateDistortionLoss
from .model import LMSFC
from .optim import net_aux_optimizer
from .parallel import BucketedGradReducer


class TrainStep:
    def __init__(self, model: LMSFC, lmbda=0.0130, lr=1e-4, aux_lr=1e-3, clip_max_norm=1.0, distributed=False,
                 bucket_mb=16.0, side_streams=True):
        self.model = model
        if model.arena is None:
            model.build_arena()
        self.arena = model.arena
        self.criterion = RateDistortionLoss(lmbda=lmbda, spec=model.spec)
        self.opt = net_aux_optimizer(model, {"net": {"type": "Adam", "lr": lr}, "aux": {"type": "Adam", "lr": aux_lr}},
                                     clip_max_norm=clip_max_norm)
        self.reducer = BucketedGradReducer(model, bucket_mb=bucket_mb) if distributed else None
        if side_streams:
            Fn.SideStreams.enable(self.arena.device)
        self.graph = None
        self.static_in = None
        self.static_out = None
        self.launches_per_step = None

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
        opt["net"].step()                       # clip_grad_norm_ + Adam, bf16 shadow in the same pass
        aux = model.aux_loss()
        aux.backward()
        opt["aux"].step()
        return {"loss": crit["loss"].detach(), "mse_loss": crit["mse_loss"].detach(),
                "bpp_loss": crit["bpp_loss"].detach(), "aux_loss": aux.detach()}

    def step(self, feats, noise_z=None, noise_y=None):
        if self.graph is not None and noise_z is None and noise_y is None:
            return self.replay(feats)
        return self._iteration(feats, noise_z, noise_y)

    # ------------------------------------------------------------------ CUDA graph
    def capture(self, feats, warmup=2):
        """Record one iteration.  `feats` fixes the shapes; its values are copied into static
        input buffers on every replay."""
        lv = self.model.spec.levels
        dev = self.arena.device
        self.static_in = {k: Fn.nhwc(feats[k], Fn.BF16).clone(memory_format=torch.channels_last) for k in lv}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):             # allocator / lazy-init warm-up outside capture
                self._iteration(self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = self._iteration(self.static_in)
        return self

    def replay(self, feats=None):
        if feats is not None:
            for k, buf in self.static_in.items():
                src = feats[k]
                if src.data_ptr() != buf.data_ptr():
                    buf.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out

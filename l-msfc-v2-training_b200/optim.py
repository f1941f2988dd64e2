"""Optimiser side of the train step: the "net" / "aux" parameter split of
compressai/optimizers/net_aux.py:39-74 and clip_grad_norm_ + Adam (examples/train.py:100-107) as
one fused launch over the flat arena (lmsfc_sumsq + lmsfc_adam_prep + lmsfc_adam_update).

`FusedAdam` IS a `torch.optim.Optimizer`, so the reference script drives it unchanged:
`ReduceLROnPlateau(optimizer, "min")` (examples/train.py:279) writes `param_groups[0]["lr"]`, which the
kernel reads from device memory (a captured CUDA graph follows the schedule), and
`optimizer.state_dict()` / `load_state_dict()` (examples/train.py:283-290, 310-321) round-trip the
Adam moments and the step counter.
"""
import torch

from . import _lib as L
from ._lib import ptr, stream


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (betas (0.9, 0.999), eps 1e-8, no weight decay / amsgrad) over a
    flat fp32 arena.  `clip_max_norm` > 0 folds torch.nn.utils.clip_grad_norm_ into the update:
    the squared-gradient sum is one reduction launch, the clip coefficient is computed on the
    device (no host sync, so the step stays capturable).

    `named_params`: the nn.Parameters living in the arena (they become `param_groups[0]["params"]`, which is what
    torch's schedulers / checkpoint code look at); `params` / `grads`: the flat arena tensors the kernels work on."""

    def __init__(self, params, grads, lr, betas=(0.9, 0.999), eps=1e-8, bf16_shadow=None, clip_max_norm=0.0,
                 zero_buffer=None, after_step=None, named_params=None):
        plist = [p for _, p in named_params] if named_params else [params]
        super().__init__(plist, dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), eps=float(eps),
                                     weight_decay=0.0, amsgrad=False))
        self.flat_params, self.grads = params, grads
        self.zero_buffer = grads if zero_buffer is None else zero_buffer    # may extend past `grads` (GDN scratch)
        self.after_step = after_step                                        # e.g. ParamArena.prepare
        self.clip_max_norm = float(clip_max_norm)
        self.m = torch.zeros_like(params)
        self.v = torch.zeros_like(params)
        self.bf16_shadow = bf16_shadow
        dev = params.device
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.scratch = torch.zeros(8, dtype=torch.float32, device=dev)   # clip, bc1, bc2, grad norm, lr, active
        self.scratch[5] = 1.0
        self.lr_dev = torch.full((1,), float(lr), dtype=torch.float32, device=dev)
        self._lr_on_device = float(lr)
        self.pending_flag = None      # device int32 [1], set by the pipelined train step (train.TrainStep)
        self.gate = None              # another FusedAdam whose "active" decision this one follows

    # ---- hyper-parameters live in param_groups[0] (what torch schedulers write)
    @property
    def lr(self):
        return float(self.param_groups[0]["lr"])

    @lr.setter
    def lr(self, value):
        self.param_groups[0]["lr"] = float(value)

    @property
    def betas(self):
        return self.param_groups[0]["betas"]

    @property
    def eps(self):
        return self.param_groups[0]["eps"]

    def sync_lr(self):
        """Mirror param_groups[0]["lr"] into device memory (a small fill; never inside a graph capture)."""
        lr = self.lr
        if lr != self._lr_on_device and not torch.cuda.is_current_stream_capturing():
            self.lr_dev.fill_(lr)
            self._lr_on_device = lr

    def zero_grad(self, set_to_none=False):
        """The gradients are views of one arena the backward kernels accumulate into: they are cleared, never dropped."""
        self.zero_buffer.zero_()

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.accumulate_norm()
        self.prep()
        self.update(0, self.flat_params.numel())
        if self.after_step is not None:
            self.after_step()
        return loss

    # ---- the same step in parts (pipelined train step: norm at the end of backward, update slices next replay)
    def accumulate_norm(self):
        """sum(g^2) of the whole arena -> self.sumsq (clip_grad_norm_'s total norm, squared); raises the pending flag."""
        if self.clip_max_norm > 0:
            self.sumsq.zero_()
            L.sumsq(ptr(self.grads), self.flat_params.numel(), ptr(self.sumsq), ptr(self.pending_flag), stream())
        elif self.pending_flag is not None:
            self.pending_flag.fill_(1)

    def prep(self, use_pending=False):
        """step += 1; clip coefficient, bias corrections and the learning rate -> device scratch.  use_pending: apply
        the update only if the device-side pending flag is raised (and lower it)."""
        self.sync_lr()
        b1, b2 = self.betas
        L.adam_prep(self.lr, ptr(self.lr_dev), b1, b2, self.clip_max_norm,
                    ptr(self.sumsq) if self.clip_max_norm > 0 else None, ptr(self.step_count), ptr(self.scratch),
                    ptr(self.pending_flag), 0 if use_pending else 1,
                    ptr(self.gate.scratch[5:]) if (use_pending and self.gate is not None) else None, stream())

    def update(self, lo, hi, zero_grad=False):
        """Element-wise Adam on arena slice [lo, hi) (multiples of 4 floats); optionally clears its gradient."""
        if hi <= lo:
            return
        sh = self.bf16_shadow[lo:hi] if self.bf16_shadow is not None else None
        b1, b2 = self.betas
        L.adam_update(ptr(self.flat_params[lo:hi]), ptr(self.m[lo:hi]), ptr(self.v[lo:hi]), ptr(self.grads[lo:hi]), ptr(sh),
                      hi - lo, b1, b2, self.eps, ptr(self.scratch), int(zero_grad), stream())

    @property
    def grad_norm(self):
        return self.scratch[3]

    # ---- checkpointing (examples/train.py:283-290, 310-321): the arena is one flat "parameter"
    def state_dict(self):
        groups = [{k: v for k, v in g.items() if k != "params"} | {"params": [0]} for g in self.param_groups]
        return {"state": {0: {"step": self.step_count.clone(), "exp_avg": self.m.clone(), "exp_avg_sq": self.v.clone()}},
                "param_groups": groups, "clip_max_norm": self.clip_max_norm}

    @torch.no_grad()
    def load_state_dict(self, sd):
        st = sd["state"][0] if "state" in sd else {"step": sd["step"], "exp_avg": sd["m"], "exp_avg_sq": sd["v"]}
        if st["exp_avg"].numel() != self.m.numel():
            raise ValueError(f"optimizer state of {st['exp_avg'].numel()} elements does not fit an arena of {self.m.numel()}")
        self.m.copy_(st["exp_avg"].reshape(-1))
        self.v.copy_(st["exp_avg_sq"].reshape(-1))
        self.step_count.copy_(torch.as_tensor(st["step"]).reshape(-1).to(torch.int32))
        for g, src in zip(self.param_groups, sd.get("param_groups", [])):
            for k, v in src.items():
                if k != "params":
                    g[k] = v
        if "lr" in sd and "param_groups" not in sd:
            self.lr = float(sd["lr"])
        self.sync_lr()


def net_aux_optimizer(model, conf=None, clip_max_norm=0.0):
    """optimizers/net_aux.py:39-74: {"net": Adam(lr), "aux": Adam(aux lr)} over the model's arenas.
    conf: {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}.
    clip_max_norm > 0 folds clip_grad_norm_ (examples/train.py:100-101) into the net optimiser's step; with the
    default 0 the caller clips, as the reference loop does."""
    conf = conf or {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}
    for k in ("net", "aux"):
        if conf[k].get("type", "Adam") != "Adam":
            raise ValueError(f"only Adam is implemented on the fused path, got {conf[k]['type']}")
    arena = model.arena or model.build_arena()
    net = FusedAdam(arena.net_params, arena.net_grads, conf["net"]["lr"], bf16_shadow=arena.net_bf16,
                    clip_max_norm=clip_max_norm, zero_buffer=arena._grads_all, after_step=arena.prepare,
                    named_params=arena.net)
    aux = FusedAdam(arena.aux_params, arena.aux_grads, conf["aux"]["lr"], named_params=arena.aux)
    return {"net": net, "aux": aux}

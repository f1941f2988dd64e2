"""Optimiser side of the train step: the "net" / "aux" parameter split of
compressai/optimizers/net_aux.py:39-74 and clip_grad_norm_ + Adam (examples/train.py:100-107) as
one fused launch over the flat arena (lmsfc_sumsq + lmsfc_adam_step).
"""
import torch

from . import _lib as L
from ._lib import ptr, stream


class FusedAdam:
    """torch.optim.Adam semantics (betas (0.9, 0.999), eps 1e-8, no weight decay / amsgrad) over a
    flat fp32 arena.  `clip_max_norm` > 0 folds torch.nn.utils.clip_grad_norm_ into the update:
    the squared-gradient sum is one reduction launch, the clip coefficient is computed on the
    device (no host sync, so the step stays capturable)."""

    def __init__(self, params, grads, lr, betas=(0.9, 0.999), eps=1e-8, bf16_shadow=None, clip_max_norm=0.0,
                 zero_buffer=None, after_step=None):
        self.params, self.grads = params, grads
        self.zero_buffer = grads if zero_buffer is None else zero_buffer    # may extend past `grads` (GDN scratch)
        self.after_step = after_step                                        # e.g. ParamArena.prepare
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.clip_max_norm = float(clip_max_norm)
        self.m = torch.zeros_like(params)
        self.v = torch.zeros_like(params)
        self.bf16_shadow = bf16_shadow
        dev = params.device
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.scratch = torch.zeros(4, dtype=torch.float32, device=dev)   # clip, bc1, bc2, grad norm

    def zero_grad(self, set_to_none=False):
        self.zero_buffer.zero_()

    def step(self):
        self.accumulate_norm()
        self.prep()
        self.update(0, self.params.numel())
        if self.after_step is not None:
            self.after_step()

    # ---- the same step in parts (pipelined train step: norm at the end of backward, update slices next replay)
    def accumulate_norm(self):
        """sum(g^2) of the whole arena -> self.sumsq (clip_grad_norm_'s total norm, squared)."""
        if self.clip_max_norm > 0:
            self.sumsq.zero_()
            L.sumsq(ptr(self.grads), self.params.numel(), ptr(self.sumsq), stream())

    def prep(self):
        """step += 1; clip coefficient and bias corrections -> device scratch."""
        L.adam_prep(self.betas[0], self.betas[1], self.clip_max_norm, ptr(self.sumsq) if self.clip_max_norm > 0 else None,
                    ptr(self.step_count), ptr(self.scratch), stream())

    def update(self, lo, hi, zero_grad=False):
        """Element-wise Adam on arena slice [lo, hi) (multiples of 4 floats); optionally clears its gradient."""
        if hi <= lo:
            return
        sh = self.bf16_shadow[lo:hi] if self.bf16_shadow is not None else None
        L.adam_update(ptr(self.params[lo:hi]), ptr(self.m[lo:hi]), ptr(self.v[lo:hi]), ptr(self.grads[lo:hi]), ptr(sh),
                      hi - lo, self.lr, self.betas[0], self.betas[1], self.eps, ptr(self.scratch), int(zero_grad), stream())

    @property
    def grad_norm(self):
        return self.scratch[3]

    def state_dict(self):
        return {"m": self.m, "v": self.v, "step": self.step_count, "lr": self.lr}

    def load_state_dict(self, sd):
        self.m.copy_(sd["m"]); self.v.copy_(sd["v"]); self.step_count.copy_(sd["step"]); self.lr = float(sd["lr"])


def net_aux_optimizer(model, conf=None, clip_max_norm=1.0):
    """optimizers/net_aux.py:39-74: {"net": Adam(lr), "aux": Adam(aux lr)} over the model's arenas.
    conf: {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}."""
    conf = conf or {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}
    for k in ("net", "aux"):
        if conf[k].get("type", "Adam") != "Adam":
            raise ValueError(f"only Adam is implemented on the fused path, got {conf[k]['type']}")
    arena = model.arena or model.build_arena()
    net = FusedAdam(arena.net_params, arena.net_grads, conf["net"]["lr"], bf16_shadow=arena.net_bf16,
                    clip_max_norm=clip_max_norm, zero_buffer=arena._grads_all, after_step=arena.prepare)
    aux = FusedAdam(arena.aux_params, arena.aux_grads, conf["aux"]["lr"])
    return {"net": net, "aux": aux}

"""Host-side mirrors of compressai.ops (bound_ops.py, parametrizers.py, ops.py).

These are parameter-sized / scalar helpers kept for interface compatibility (state_dict keys,
init values); the activation-sized work they imply is fused into the CUDA kernels
(lmsfc_reparam_forward/backward, the LowerBound clauses inside the likelihood kernels).
"""
import torch
import torch.nn as nn

from . import _lib as L
from ._lib import ptr, stream


class _LowerBoundFn(torch.autograd.Function):
    """compressai/ops/bound_ops.py:36-57: fwd max(x, b); bwd g * [(x >= b) | (g < 0)]."""

    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, g):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (g < 0)) * g, None


class LowerBound(nn.Module):
    """compressai/ops/bound_ops.py:59-80 (same buffer name: `bound`)."""

    def __init__(self, bound: float):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))

    def forward(self, x):
        return _LowerBoundFn.apply(x, self.bound)


class NonNegativeParametrizer(nn.Module):
    """compressai/ops/parametrizers.py:38-64 (buffers `pedestal`, `lower_bound.bound`)."""

    def __init__(self, minimum: float = 0, reparam_offset: float = 2 ** -18):
        super().__init__()
        self.minimum = float(minimum)
        self.reparam_offset = float(reparam_offset)
        pedestal = self.reparam_offset ** 2
        self.register_buffer("pedestal", torch.Tensor([pedestal]))
        self.bound_value = (self.minimum + self.reparam_offset ** 2) ** 0.5
        self.lower_bound = LowerBound(self.bound_value)

    def init(self, x):
        return torch.sqrt(torch.max(x + self.pedestal, self.pedestal))

    def forward(self, x):
        """Parameter-sized evaluation through the CUDA kernel (used by tests / introspection;
        GDN.forward fuses this into its own launch sequence)."""
        if not x.is_cuda:
            raise RuntimeError("NonNegativeParametrizer runs on CUDA tensors only (no CPU fallback)")
        xc = x.detach().contiguous().float()
        out = torch.empty_like(xc)
        L.reparam_forward(ptr(xc), xc.numel(), self.bound_value, self.reparam_offset ** 2, ptr(out), None, stream())
        return out


def quantize_ste(x):
    """compressai/ops/ops.py:66-80: round with a straight-through gradient."""
    return (torch.round(x) - x).detach() + x

"""Drop-in GDN / conv / deconv modules (interfaces of compressai.layers.GDN and
compressai.models.utils.conv/deconv) running on the tcgen05 implicit-GEMM kernels.

state_dict keys, shapes and init values equal the reference's:
  GDN: beta [C], gamma [C,C], beta_reparam.pedestal, beta_reparam.lower_bound.bound,
       gamma_reparam.pedestal, gamma_reparam.lower_bound.bound   (layers/gdn.py:54-75)
  conv: weight [Cout,Cin,k,k], bias [Cout]; deconv: weight [Cin,Cout,k,k], bias [Cout]
The conv weights are *stored* in GEMM layout ([D0][kh][kw][D1], tap-major, channel-contiguous) and
exposed as permuted views, so indices and shapes are the reference's while the tensor cores read
them without a transpose.
"""
import math

import torch
import torch.nn as nn

from . import functional as Fn
from .ops import NonNegativeParametrizer


class GDN(nn.Module):
    """y = x * rsqrt(beta' + gamma' . x^2) (inverse=True: x * sqrt(...)); layers/gdn.py:41-92."""

    def __init__(self, in_channels: int, inverse: bool = False, beta_min: float = 1e-6, gamma_init: float = 0.1):
        super().__init__()
        beta_min = float(beta_min)
        gamma_init = float(gamma_init)
        self.inverse = bool(inverse)
        self.in_channels = int(in_channels)
        self.beta_reparam = NonNegativeParametrizer(minimum=beta_min)
        self.beta = nn.Parameter(self.beta_reparam.init(torch.ones(in_channels)))
        self.gamma_reparam = NonNegativeParametrizer()
        self.gamma = nn.Parameter(self.gamma_reparam.init(gamma_init * torch.eye(in_channels)))
        self._state = Fn.GDNState(in_channels, inverse, self.beta_reparam.bound_value, self.gamma_reparam.bound_value)

    def forward(self, x):
        if x.dim() != 4 or x.shape[1] != self.in_channels:
            raise ValueError(f"GDN({self.in_channels}) got input of shape {tuple(x.shape)}")
        return Fn.gdn_apply(self._state, x, self.beta, self.gamma)


class _ConvBase(nn.Module):
    kind = 0

    def __init__(self, in_channels, out_channels, kernel_size=5, stride=2, concat_channels=0):
        """`concat_channels`: trailing input channels that arrive as a second tensor (the FPN level
        fused at this scale) -- the kernel reads both sources, no torch.cat is materialised."""
        super().__init__()
        self.in_channels, self.out_channels = int(in_channels), int(out_channels)
        self.kernel_size, self.stride = int(kernel_size), int(stride)
        self.concat_channels = int(concat_channels)
        c0 = self.in_channels - self.concat_channels
        self._state = Fn.ConvState(self.kind, self.kernel_size, self.stride, c0, self.concat_channels, self.out_channels)
        k = self.kernel_size
        d0, d1 = self._state.D0, self._state.D1
        # torch default init (kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(fan_in), 1/sqrt(fan_in))) on the
        # reference-shaped tensor, then re-laid out; fan_in follows torch's rule for each module type
        w = torch.empty(d0, d1, k, k)
        nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        fan_in = d1 * k * k
        storage = w.permute(0, 2, 3, 1).contiguous()            # [D0][kh][kw][D1]
        self.weight = nn.Parameter(storage.permute(0, 3, 1, 2))  # logical [D0][D1][kh][kw]
        bound = 1 / math.sqrt(fan_in)
        self.bias = nn.Parameter(torch.empty(self.out_channels).uniform_(-bound, bound))

    def _apply(self, fn, recurse=True):
        # nn.Module._apply may re-create parameters; drop the bf16 shadows so they are rebuilt
        out = super()._apply(fn, recurse)
        self._state.w_direct = None
        self._state._version = -1
        return out

    def forward(self, x, x_cat=None, relu=False, out_hw=None, latent_head=False, link=None, side=False):
        """relu: False / True / "leaky" (LeakyReLU(0.01)).  latent_head: True -> (fp32 y, bf16 |y|), "plain" -> (fp32 y,
        bf16 y).  `link` / `side`: the GDNLink of the GDN that produced `x` (see functional.GDNLink) -- lets this
        layer's data-gradient kernel run that GDN's backward contraction in its epilogue."""
        if (x_cat is None) != (self.concat_channels == 0):
            raise ValueError("concat source mismatch")
        mode = 2 if latent_head == "plain" else (1 if latent_head else 0)
        return Fn.conv_apply(self._state, x, x_cat, self.weight, self.bias, relu=relu, mode=mode,
                             out_hw=out_hw, link=link, side=side)

    def forward_gdn(self, gdn, x, x_cat=None, out_hw=None, link=None):
        """gdn(self(x)) as one autograd node, the GDN contraction fused into this layer's kernel where the tiling
        allows it (functional._ConvGDNFn).  -> (y, GDNLink for y's consumers)."""
        if (x_cat is None) != (self.concat_channels == 0):
            raise ValueError("concat source mismatch")
        return Fn.conv_gdn_apply(self._state, gdn._state, x, x_cat, self.weight, self.bias, gdn.beta, gdn.gamma,
                                 out_hw=out_hw, in_link=link)


class Conv(_ConvBase):
    """nn.Conv2d(k, stride, padding=k//2) -- models/utils.py:128-135."""
    kind = 0


class Deconv(_ConvBase):
    """nn.ConvTranspose2d(k, stride, padding=k//2, output_padding=stride-1), output cropped to
    `out_hw` -- models/utils.py:138-146."""
    kind = 1


def conv(in_channels, out_channels, kernel_size=5, stride=2, concat_channels=0):
    return Conv(in_channels, out_channels, kernel_size, stride, concat_channels)


def deconv(in_channels, out_channels, kernel_size=5, stride=2):
    return Deconv(in_channels, out_channels, kernel_size, stride)

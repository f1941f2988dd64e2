"""The L-MSFC codec model (topology frozen in spec.py) on the B200 kernels.

Contract kept from compressai.models.base.CompressionModel (models/base.py:62-144) and
models/google.py:292-303: `forward(x) -> {"x_hat": ..., "likelihoods": {...}}`, `aux_loss()`,
`update(scale_table=None, force=False)`, `load_state_dict` resizing the CDF buffers first.
Here `x` and `x_hat` are dicts of the pyramid levels (p2..p5, [B,F,H,W]).

ParamArena packs every fp32 parameter (and its gradient) into one flat allocation so that
clip+Adam is one launch, the bf16 GEMM operands of all conv weights are one contiguous block
written by the same launch (each is read K-major or MN-major by the kernels: no transposed copies), and gradient all-reduce buckets are plain slices.
"""
import math

import torch
import torch.nn as nn

from . import _lib as L
from . import functional as Fn
from ._lib import ptr, stream
from .entropy_models import EntropyBottleneck, GaussianConditional
from .layers import GDN, Conv, Deconv, _ConvBase
from .spec import DEFAULT_SPEC, LMSFCSpec

SCALES_MIN, SCALES_MAX, SCALES_LEVELS = 0.11, 256, 64


def get_scale_table(min=SCALES_MIN, max=SCALES_MAX, levels=SCALES_LEVELS):
    """compressai/models/base.py:57-59."""
    return torch.exp(torch.linspace(math.log(min), math.log(max), levels))


class LMSFC(nn.Module):
    def __init__(self, spec: LMSFCSpec = DEFAULT_SPEC):
        super().__init__()
        self.spec = spec
        Fc, N, M = spec.F, spec.N, spec.M
        g_a = nn.Module()
        g_a.conv2, g_a.gdn2 = Conv(Fc, N), GDN(N)
        g_a.conv3, g_a.gdn3 = Conv(N + Fc, N, concat_channels=Fc), GDN(N)
        g_a.conv4, g_a.gdn4 = Conv(N + Fc, N, concat_channels=Fc), GDN(N)
        g_a.conv5 = Conv(N + Fc, M, concat_channels=Fc)
        self.g_a = g_a
        # keys "0","2","4" mirror the nn.Sequential(conv, ReLU, conv, ReLU, conv) of google.py:267-282
        self.h_a = nn.ModuleDict({"0": Conv(M, N, 3, 1), "2": Conv(N, N), "4": Conv(N, N)})
        self.h_s = nn.ModuleDict({"0": Deconv(N, N), "2": Deconv(N, N), "4": Conv(N, M, 3, 1)})
        self.entropy_bottleneck = EntropyBottleneck(N)
        self.gaussian_conditional = GaussianConditional(None, scale_bound=spec.scale_bound,
                                                        likelihood_bound=spec.likelihood_bound)
        g_s = nn.Module()
        g_s.up5, g_s.igdn5, g_s.head5 = Deconv(M, N), GDN(N, inverse=True), Conv(N, Fc, 3, 1)
        g_s.up4, g_s.igdn4, g_s.head4 = Deconv(N, N), GDN(N, inverse=True), Conv(N, Fc, 3, 1)
        g_s.up3, g_s.igdn3, g_s.head3 = Deconv(N, N), GDN(N, inverse=True), Conv(N, Fc, 3, 1)
        g_s.up2 = Deconv(N, Fc)
        self.g_s = g_s
        self.arena = None

    # ------------------------------------------------------------------ forward
    def forward(self, feats, noise_z=None, noise_y=None):
        """feats: {"p2": [B,F,H,W], ...}.  noise_z / noise_y (optional, shaped like z / y) inject
        the U(-.5,.5) draws for parity tests; RNG order is EB then GC (google.py:295-297)."""
        lv = self.spec.levels
        p2, p3, p4, p5 = (Fn.nhwc(feats[k], Fn.BF16) for k in lv)
        a, s = self.g_a, self.g_s
        e2 = a.gdn2(a.conv2(p2))
        e3 = a.gdn3(a.conv3(e2, p3))
        e4 = a.gdn4(a.conv4(e3, p4))
        y, y_abs = a.conv5(e4, p5, latent_head=True)
        z1 = self.h_a["0"](y_abs, relu=True)
        z2 = self.h_a["2"](z1, relu=True)
        z, _ = self.h_a["4"](z2, latent_head=True)
        z_hat, z_lik = self.entropy_bottleneck(z, noise=noise_z)
        t = self.h_s["0"](z_hat.to(Fn.BF16), relu=True, out_hw=z2.shape[2:])
        t = self.h_s["2"](t, relu=True, out_hw=y.shape[2:])
        scales, _ = self.h_s["4"](t, relu=True, latent_head=True)
        y_hat, y_lik = self.gaussian_conditional(y, scales, noise=noise_y)
        d5 = s.igdn5(s.up5(y_hat.to(Fn.BF16), out_hw=p5.shape[2:]))
        p5h = s.head5(d5)
        d4 = s.igdn4(s.up4(d5, out_hw=p4.shape[2:]))
        p4h = s.head4(d4)
        d3 = s.igdn3(s.up3(d4, out_hw=p3.shape[2:]))
        p3h = s.head3(d3)
        p2h = s.up2(d3, out_hw=p2.shape[2:])
        return {"x_hat": dict(zip(lv, (p2h, p3h, p4h, p5h))),
                "likelihoods": {"y": y_lik, "z": z_lik},
                "latents": {"y": y, "z": z, "y_hat": y_hat, "z_hat": z_hat, "scales": scales},
                "inputs_nhwc": dict(zip(lv, (p2, p3, p4, p5)))}

    # ------------------------------------------------------------------ CompressionModel contract
    def aux_loss(self):
        """models/base.py:115-144: sum of the EntropyBottleneck aux losses."""
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))

    def update(self, scale_table=None, force=False):
        """models/base.py:90-113."""
        if scale_table is None:
            scale_table = get_scale_table()
        updated = self.gaussian_conditional.update_scale_table(scale_table, force=force)
        for m in self.modules():
            if isinstance(m, EntropyBottleneck):
                updated |= m.update(force=force)
        return updated

    def load_state_dict(self, state_dict, strict=True):
        """models/base.py:67-88: CDF buffers are empty until update(); resize them to the checkpoint."""
        for name, module in self.named_modules():
            if not isinstance(module, (EntropyBottleneck, GaussianConditional)):
                continue
            buffers = ["_quantized_cdf", "_offset", "_cdf_length"]
            if isinstance(module, GaussianConditional):
                buffers.append("scale_table")
            for b in buffers:
                key = f"{name}.{b}"
                if key in state_dict:
                    src = state_dict[key]
                    cur = getattr(module, b)
                    if cur.shape != src.shape:
                        setattr(module, b, torch.empty(src.shape, dtype=cur.dtype, device=cur.device))
        out = super().load_state_dict(state_dict, strict=strict)
        self.refresh_shadows()
        return out

    # ------------------------------------------------------------------ arena / shadows
    def conv_layers(self):
        return [m for m in self.modules() if isinstance(m, _ConvBase)]

    def build_arena(self):
        """Move all parameters (and .grad) into flat fp32 arenas; idempotent per device."""
        self.arena = ParamArena(self)
        return self.arena

    def refresh_shadows(self, cast=True):
        """Rebuild the bf16 GEMM operands after the fp32 weights changed (optimizer step / load)."""
        if self.arena is not None:
            self.arena.refresh_shadows(cast=cast)
        else:
            for m in self.conv_layers():
                if m.weight.is_cuda:
                    m._state.refresh(m.weight, force=True)


def _round_up(n, k):
    return (n + k - 1) // k * k


class ParamArena:
    """Flat fp32 storage for parameters / gradients (+ one bf16 shadow block).

    "net" holds every parameter except `*.quantiles`, "aux" holds those -- the split of
    compressai/optimizers/net_aux.py:46-57.  Every parameter keeps its logical shape and strides
    (conv weights stay GEMM-layout permuted views), only its storage moves."""

    ALIGN = 64   # floats: 256-byte aligned slices (TMA bases, float4 kernels)

    def __init__(self, model):
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        if not named or not named[0][1].is_cuda:
            raise RuntimeError("build_arena(): move the model to a CUDA device first (no CPU fallback)")
        dev = named[0][1].device
        self.device = dev
        self.net = [(n, p) for n, p in named if not n.endswith(".quantiles")]
        self.aux = [(n, p) for n, p in named if n.endswith(".quantiles")]
        self.offsets = {}

        def pack(group):
            off = 0
            for n, p in group:
                self.offsets[n] = off
                off += _round_up(p.numel(), self.ALIGN)
            flat = torch.zeros(max(off, self.ALIGN), dtype=torch.float32, device=dev)
            grad = torch.zeros_like(flat)
            for n, p in group:
                o = self.offsets[n]
                view = flat.as_strided(p.shape, p.stride(), o)
                view.copy_(p.data)
                p.data = view
                p.grad = grad.as_strided(p.shape, p.stride(), o)
            return flat, grad

        with torch.no_grad():
            self.net_params, self.net_grads = pack(self.net)
            self.aux_params, self.aux_grads = pack(self.aux)
        self.net_bf16 = torch.empty(self.net_params.numel(), dtype=torch.bfloat16, device=dev)
        self.model = model
        for m in model.conv_layers():
            name = next(n for n, p in self.net if p is m.weight)
            o = self.offsets[name]
            st = m._state
            st.w_direct = self.net_bf16[o:o + m.weight.numel()].view(st.D0, st.T, st.D1)
        self.refresh_shadows(cast=True)

    def names(self, group="net"):
        return [n for n, _ in (self.net if group == "net" else self.aux)]

    def slice_of(self, name, what="grad"):
        base = {"grad": self.net_grads, "param": self.net_params}[what]
        p = dict(self.net)[name]
        o = self.offsets[name]
        return base[o:o + p.numel()]

    def refresh_shadows(self, cast=True):
        """bf16 GEMM operands of all conv weights = one cast of the flat arena (Adam writes the same
        block in its own pass, so the train step never calls this)."""
        if cast:
            L.cast_f32_bf16(ptr(self.net_params), ptr(self.net_bf16), self.net_params.numel(), stream())
        for m in self.model.conv_layers():
            m._state._version = m.weight._version

    def zero_grad(self):
        self.net_grads.zero_()
        self.aux_grads.zero_()

"""Oracle (test infrastructure, see oracle/__init__.py): the frozen L-MSFC graph
(spec.py of the product package) assembled from the restated reference maths in
oracle/ops.py, plus the reference train step (compressai/examples/train.py:
85-118) and optimiser split (optimizers/net_aux.py:39-74).  fp32, any device.

Parameter names equal the product model's so one state_dict serves both.
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops as O


def conv(cin, cout, k=5, s=2):
    """models/utils.py:128-135."""
    return nn.Conv2d(cin, cout, k, stride=s, padding=k // 2)


def deconv(cin, cout, k=5, s=2):
    """models/utils.py:138-146."""
    return nn.ConvTranspose2d(cin, cout, k, stride=s, output_padding=s - 1, padding=k // 2)


class GDNParams(nn.Module):
    """layers/gdn.py:54-75 (init: beta=1, gamma=0.1*I through the reparam init)."""

    def __init__(self, C, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse, self.beta_min = bool(inverse), float(beta_min)
        self.beta = nn.Parameter(O.nonneg_init(torch.ones(C)))
        self.gamma = nn.Parameter(O.nonneg_init(gamma_init * torch.eye(C)))

    def forward(self, x):
        return O.gdn(x, self.beta, self.gamma, self.inverse, self.beta_min)


class EBParams(nn.Module):
    """entropy_models.py:342-383."""

    def __init__(self, C, init_scale=10.0, filters=(3, 3, 3, 3), tail_mass=1e-9):
        super().__init__()
        self.channels, self.filters = C, tuple(filters)
        f = (1,) + self.filters + (1,)
        scale = init_scale ** (1 / (len(self.filters) + 1))
        for i in range(len(self.filters) + 1):
            init = np.log(np.expm1(1 / scale / f[i + 1]))
            self.register_parameter(f"_matrix{i}", nn.Parameter(torch.full((C, f[i + 1], f[i]), float(init))))
            self.register_parameter(f"_bias{i}", nn.Parameter(torch.empty(C, f[i + 1], 1).uniform_(-0.5, 0.5)))
            if i < len(self.filters):
                self.register_parameter(f"_factor{i}", nn.Parameter(torch.zeros(C, f[i + 1], 1)))
        self.quantiles = nn.Parameter(torch.tensor([-init_scale, 0.0, init_scale]).repeat(C, 1, 1))
        t = np.log(2 / tail_mass - 1)
        self.register_buffer("target", torch.tensor([-t, 0.0, t], dtype=torch.float32))

    def mbf(self):
        n = len(self.filters) + 1
        return ([getattr(self, f"_matrix{i}") for i in range(n)],
                [getattr(self, f"_bias{i}") for i in range(n)],
                [getattr(self, f"_factor{i}") for i in range(n - 1)])

    def forward(self, x, training=True, noise=None):
        m, b, f = self.mbf()
        return O.eb_forward(x, m, b, f, self.quantiles, training, noise)

    def loss(self):
        m, b, f = self.mbf()
        return O.eb_aux_loss(m, b, f, self.quantiles, self.target)


class OracleLMSFC(nn.Module):
    """The graph spec.py freezes (any number of levels / per-level channels), from the restated reference modules."""

    def __init__(self, spec):
        super().__init__()
        self.spec = spec
        N, M, ch = spec.N, spec.M, spec.channels
        n = len(spec.levels)
        self._tags = [spec.tag(i) for i in range(n)]
        g_a = nn.Module()
        for i, t in enumerate(self._tags):
            last = i == n - 1
            setattr(g_a, f"conv{t}", conv(ch[i] + (N if i else 0), M if last else N))
            if not last:
                setattr(g_a, f"gdn{t}", GDNParams(N))
        self.g_a = g_a
        self.mean_scale = getattr(spec, "hyperprior", "scale") == "mean-scale"
        if self.mean_scale:      # models/google.py:377-391; middle width from the spec (reference: M * 3 // 2)
            mid = spec.mid
            self.h_a = nn.Sequential(conv(M, N, 3, 1), nn.LeakyReLU(), conv(N, N), nn.LeakyReLU(), conv(N, N))
            self.h_s = nn.Sequential(deconv(N, M), nn.LeakyReLU(), deconv(M, mid), nn.LeakyReLU(), conv(mid, 2 * M, 3, 1),
                                     nn.Identity())
        else:                    # models/google.py:267-282
            self.h_a = nn.Sequential(conv(M, N, 3, 1), nn.ReLU(), conv(N, N), nn.ReLU(), conv(N, N))
            self.h_s = nn.Sequential(deconv(N, N), nn.ReLU(), deconv(N, N), nn.ReLU(), conv(N, M, 3, 1), nn.ReLU())
        self.entropy_bottleneck = EBParams(N)
        g_s = nn.Module()
        for i in range(n - 1, 0, -1):
            t = self._tags[i]
            setattr(g_s, f"up{t}", deconv(M if i == n - 1 else N, N))
            setattr(g_s, f"igdn{t}", GDNParams(N, True))
            setattr(g_s, f"head{t}", conv(N, ch[i], 3, 1))
        setattr(g_s, f"up{self._tags[0]}", deconv(N, ch[0]))
        self.g_s = g_s

    @staticmethod
    def _crop(t, ref):
        return t[..., : ref.shape[-2], : ref.shape[-1]]

    def _analysis(self, xs):
        a, tags = self.g_a, self._tags
        e = getattr(a, f"gdn{tags[0]}")(getattr(a, f"conv{tags[0]}")(xs[0]))
        for i in range(1, len(tags) - 1):
            e = getattr(a, f"gdn{tags[i]}")(getattr(a, f"conv{tags[i]}")(torch.cat([e, xs[i]], 1)))
        return getattr(a, f"conv{tags[-1]}")(torch.cat([e, xs[-1]], 1))

    def _synthesis(self, y_hat, level_hw):
        s, tags = self.g_s, self._tags
        n = len(tags)
        crop = lambda t, hw: t[..., : hw[0], : hw[1]]
        outs = [None] * n
        d = y_hat
        for i in range(n - 1, 0, -1):
            t = tags[i]
            d = getattr(s, f"igdn{t}")(crop(getattr(s, f"up{t}")(d), level_hw[i]))
            outs[i] = getattr(s, f"head{t}")(d)
        outs[0] = crop(getattr(s, f"up{tags[0]}")(d), level_hw[0])
        return outs

    def forward(self, feats, noise_z=None, noise_y=None, training=True):
        """feats: dict of the pyramid levels, [B,F_l,H,W] fp32.  noise_z in EB's [C,1,B*h*w]
        layout, noise_y in [B,M,h,w]; drawn here (EB first) when None."""
        lv = self.spec.levels
        xs = [feats[k] for k in lv]
        y = self._analysis(xs)
        ha = self.h_a
        z1 = ha[2](ha[1](ha[0](y if self.mean_scale else torch.abs(y))))     # google.py:395 | :294
        z = ha[4](ha[3](z1))
        if training and noise_z is None:
            noise_z = torch.empty(z.shape[1], 1, z.numel() // z.shape[1], device=z.device).uniform_(-0.5, 0.5)
        z_hat, z_lik = self.entropy_bottleneck(z, training, noise_z)
        hs = self.h_s  # every deconv output is cropped to its target size (spec.py)
        t = hs[1](self._crop(hs[0](z_hat), z1))
        t = hs[3](self._crop(hs[2](t), y))
        scales = hs[5](hs[4](t))
        means = None
        if self.mean_scale:
            scales, means = scales.chunk(2, 1)                                  # google.py:397-398
        if training and noise_y is None:
            noise_y = torch.empty_like(y).uniform_(-0.5, 0.5)
        y_hat, y_lik = O.gc_forward(y, scales, means, training, noise_y,
                                    self.spec.scale_bound, self.spec.likelihood_bound)
        outs = self._synthesis(y_hat, [x.shape[-2:] for x in xs])
        return {"x_hat": dict(zip(lv, outs)),
                "likelihoods": {"y": y_lik, "z": z_lik},
                "latents": {"y": y, "z": z, "y_hat": y_hat, "z_hat": z_hat, "scales": scales, "means": means}}


    # ------------------------------------------------------------------ entropy-coded path (models/google.py:313-333)
    def update(self):
        """models/base.py:90-113: integer CDF tables of both entropy models (entropy_models.py:389-429, 625-648)."""
        from . import cdf as OC
        m, b, f = self.entropy_bottleneck.mbf()
        self._eb_tabs = OC.eb_update([t.detach() for t in m], [t.detach() for t in b], [t.detach() for t in f],
                                     self.entropy_bottleneck.quantiles.detach())
        self._scale_table = OC.get_scale_table()
        self._gc_tabs = OC.gc_update(self._scale_table)

    def build_indexes(self, scales):
        """entropy_models.py:684-689."""
        scales = torch.clamp(scales, min=self.spec.scale_bound)
        indexes = scales.new_full(scales.size(), len(self._scale_table) - 1).int()
        for s in self._scale_table[:-1]:
            indexes -= (scales <= s).int()
        return indexes

    def _scales_from(self, z_hat, mid_hw, y_hw):
        hs = self.h_s
        t = hs[1](hs[0](z_hat)[..., : mid_hw[0], : mid_hw[1]])
        t = hs[3](hs[2](t)[..., : y_hw[0], : y_hw[1]])
        p = hs[5](hs[4](t))
        return tuple(p.chunk(2, 1)) if self.mean_scale else (p, None)

    @torch.no_grad()
    def compress(self, feats):
        """ScaleHyperprior.compress (models/google.py:313-324) with the oracle's coder (oracle/rans.py)."""
        from . import rans as OR
        y = self._analysis([feats[k] for k in self.spec.levels])
        ha = self.h_a
        z1 = ha[2](ha[1](ha[0](y if self.mean_scale else torch.abs(y))))
        z = ha[4](ha[3](z1))
        med = self.entropy_bottleneck.quantiles[:, 0, 1].reshape(1, -1, 1, 1)
        cdf, off, ln = self._eb_tabs
        z_sym = torch.round(z - med).int().numpy()
        z_idx = np.broadcast_to(np.arange(z.shape[1], dtype=np.int32).reshape(1, -1, 1, 1), z_sym.shape)
        z_strings = OR.compress(z_sym, z_idx, cdf, ln, off)
        z_hat = torch.from_numpy(OR.decompress(z_strings, z_idx, cdf, ln, off)).float() + med
        scales, means = self._scales_from(z_hat, z1.shape[-2:], y.shape[-2:])
        indexes = self.build_indexes(scales)
        cdf, off, ln = self._gc_tabs
        y_sym = torch.round(y - means if means is not None else y).int()          # entropy_models.py:170-180, "symbols"
        y_strings = OR.compress(y_sym.numpy(), indexes.numpy(), cdf, ln, off)
        return {"strings": [y_strings, z_strings], "shape": tuple(z.shape[-2:]), "y_hw": tuple(y.shape[-2:]),
                "mid_hw": tuple(z1.shape[-2:]), "level_hw": [tuple(feats[k].shape[-2:]) for k in self.spec.levels]}

    @torch.no_grad()
    def decompress(self, strings, shape, y_hw, mid_hw, level_hw):
        """ScaleHyperprior.decompress (models/google.py:326-333)."""
        from . import rans as OR
        B, N = len(strings[1]), self.entropy_bottleneck.channels
        med = self.entropy_bottleneck.quantiles[:, 0, 1].reshape(1, -1, 1, 1)
        cdf, off, ln = self._eb_tabs
        z_idx = np.broadcast_to(np.arange(N, dtype=np.int32).reshape(1, -1, 1, 1), (B, N, *shape))
        z_hat = torch.from_numpy(OR.decompress(strings[1], z_idx, cdf, ln, off)).float() + med
        scales, means = self._scales_from(z_hat, mid_hw, y_hw)
        indexes = self.build_indexes(scales)
        cdf, off, ln = self._gc_tabs
        y_hat = torch.from_numpy(OR.decompress(strings[0], indexes.numpy(), cdf, ln, off)).float()
        if means is not None:
            y_hat = y_hat + means                                                 # dequantize, entropy_models.py:182-190
        return {"x_hat": dict(zip(self.spec.levels, self._synthesis(y_hat, level_hw)))}

    def aux_loss(self):
        """models/base.py:115-144."""
        return self.entropy_bottleneck.loss()

    def criterion(self, out, feats, lmbda):
        lv = self.spec.levels
        B, _, H2, W2 = feats[lv[0]].shape
        return O.rd_loss(list(out["likelihoods"].values()),
                         [out["x_hat"][k] for k in lv], [feats[k] for k in lv],
                         lmbda, self.spec.num_pixels(B, H2, W2), self.spec.level_weights)


def configure_optimizers(net, lr=1e-4, aux_lr=1e-3):
    """examples/train.py:74-82 + optimizers/net_aux.py:39-74 (Adam/Adam)."""
    named = dict(net.named_parameters())
    net_names = sorted(n for n in named if not n.endswith(".quantiles"))
    aux_names = sorted(n for n in named if n.endswith(".quantiles"))
    return (torch.optim.Adam((named[n] for n in net_names), lr=lr),
            torch.optim.Adam((named[n] for n in aux_names), lr=aux_lr))


def train_step(net, opt, aux_opt, feats, lmbda, clip_max_norm=1.0, noise_z=None, noise_y=None):
    """examples/train.py:94-107, one iteration."""
    opt.zero_grad()
    aux_opt.zero_grad()
    out = net(feats, noise_z, noise_y, training=True)
    crit = net.criterion(out, feats, lmbda)
    crit["loss"].backward()
    if clip_max_norm > 0:
        torch.nn.utils.clip_grad_norm_(net.parameters(), clip_max_norm)
    opt.step()
    aux = net.aux_loss()
    aux.backward()
    aux_opt.step()
    return {"loss": crit["loss"].detach(), "mse_loss": crit["mse_loss"].detach(),
            "bpp_loss": crit["bpp_loss"].detach(), "aux_loss": aux.detach()}


def synth_features(spec, B, H2, W2, seed=1234, device="cpu", dtype=torch.float32):
    """FPN-like features relu(N(0,1))*s_l + N(0,0.1), clipped to
    the dataset range [-26.43, 28.40] (compressai_vision/codecs/utils.py:30-39)."""
    g = torch.Generator().manual_seed(seed)
    feats = {}
    for i, (k, (h, w)) in enumerate(zip(spec.levels, spec.level_shapes(H2, W2))):
        C = spec.channels[i]
        a = torch.relu(torch.randn(B, C, h, w, generator=g)) * (1.0 + 0.5 * i)
        a = a + 0.1 * torch.randn(B, C, h, w, generator=g)
        feats[k] = a.clamp_(-26.43, 28.40).to(device=device, dtype=dtype)
    return feats

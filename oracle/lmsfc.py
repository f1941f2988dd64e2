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
    def __init__(self, spec):
        super().__init__()
        self.spec = spec
        Fc, N, M = spec.F, spec.N, spec.M
        g_a = nn.Module()
        g_a.conv2, g_a.gdn2 = conv(Fc, N), GDNParams(N)
        g_a.conv3, g_a.gdn3 = conv(N + Fc, N), GDNParams(N)
        g_a.conv4, g_a.gdn4 = conv(N + Fc, N), GDNParams(N)
        g_a.conv5 = conv(N + Fc, M)
        self.g_a = g_a
        self.h_a = nn.Sequential(conv(M, N, 3, 1), nn.ReLU(), conv(N, N), nn.ReLU(), conv(N, N))
        self.h_s = nn.Sequential(deconv(N, N), nn.ReLU(), deconv(N, N), nn.ReLU(), conv(N, M, 3, 1), nn.ReLU())
        self.entropy_bottleneck = EBParams(N)
        g_s = nn.Module()
        g_s.up5, g_s.igdn5, g_s.head5 = deconv(M, N), GDNParams(N, True), conv(N, Fc, 3, 1)
        g_s.up4, g_s.igdn4, g_s.head4 = deconv(N, N), GDNParams(N, True), conv(N, Fc, 3, 1)
        g_s.up3, g_s.igdn3, g_s.head3 = deconv(N, N), GDNParams(N, True), conv(N, Fc, 3, 1)
        g_s.up2 = deconv(N, Fc)
        self.g_s = g_s

    @staticmethod
    def _crop(t, ref):
        return t[..., : ref.shape[-2], : ref.shape[-1]]

    def forward(self, feats, noise_z=None, noise_y=None, training=True):
        """feats: dict p2..p5 of [B,F,H,W] fp32.  noise_z in EB's [C,1,B*h*w]
        layout, noise_y in [B,M,h,w]; drawn here (EB first) when None."""
        p2, p3, p4, p5 = (feats[k] for k in self.spec.levels)
        a = self.g_a
        e2 = a.gdn2(a.conv2(p2))
        e3 = a.gdn3(a.conv3(torch.cat([e2, p3], 1)))
        e4 = a.gdn4(a.conv4(torch.cat([e3, p4], 1)))
        y = a.conv5(torch.cat([e4, p5], 1))
        ha = self.h_a
        z1 = ha[2](ha[1](ha[0](torch.abs(y))))
        z = ha[4](ha[3](z1))
        if training and noise_z is None:
            noise_z = torch.empty(z.shape[1], 1, z.numel() // z.shape[1], device=z.device).uniform_(-0.5, 0.5)
        z_hat, z_lik = self.entropy_bottleneck(z, training, noise_z)
        hs = self.h_s  # every deconv output is cropped to its target size (spec.py)
        t = hs[1](self._crop(hs[0](z_hat), z1))
        t = hs[3](self._crop(hs[2](t), y))
        scales = hs[5](hs[4](t))
        if training and noise_y is None:
            noise_y = torch.empty_like(y).uniform_(-0.5, 0.5)
        y_hat, y_lik = O.gc_forward(y, scales, None, training, noise_y,
                                    self.spec.scale_bound, self.spec.likelihood_bound)
        s = self.g_s
        d5 = s.igdn5(self._crop(s.up5(y_hat), p5)); p5h = s.head5(d5)
        d4 = s.igdn4(self._crop(s.up4(d5), p4)); p4h = s.head4(d4)
        d3 = s.igdn3(self._crop(s.up3(d4), p3)); p3h = s.head3(d3)
        p2h = self._crop(s.up2(d3), p2)
        return {"x_hat": {"p2": p2h, "p3": p3h, "p4": p4h, "p5": p5h},
                "likelihoods": {"y": y_lik, "z": z_lik},
                "latents": {"y": y, "z": z, "y_hat": y_hat, "z_hat": z_hat, "scales": scales}}


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
        return hs[5](hs[4](t))

    @torch.no_grad()
    def compress(self, feats):
        """ScaleHyperprior.compress (models/google.py:313-324) with the oracle's coder (oracle/rans.py)."""
        from . import rans as OR
        p2, p3, p4, p5 = (feats[k] for k in self.spec.levels)
        a = self.g_a
        e2 = a.gdn2(a.conv2(p2))
        e3 = a.gdn3(a.conv3(torch.cat([e2, p3], 1)))
        e4 = a.gdn4(a.conv4(torch.cat([e3, p4], 1)))
        y = a.conv5(torch.cat([e4, p5], 1))
        ha = self.h_a
        z1 = ha[2](ha[1](ha[0](torch.abs(y))))
        z = ha[4](ha[3](z1))
        med = self.entropy_bottleneck.quantiles[:, 0, 1].reshape(1, -1, 1, 1)
        cdf, off, ln = self._eb_tabs
        z_sym = torch.round(z - med).int().numpy()
        z_idx = np.broadcast_to(np.arange(z.shape[1], dtype=np.int32).reshape(1, -1, 1, 1), z_sym.shape)
        z_strings = OR.compress(z_sym, z_idx, cdf, ln, off)
        z_hat = torch.from_numpy(OR.decompress(z_strings, z_idx, cdf, ln, off)).float() + med
        scales = self._scales_from(z_hat, z1.shape[-2:], y.shape[-2:])
        indexes = self.build_indexes(scales)
        cdf, off, ln = self._gc_tabs
        y_strings = OR.compress(torch.round(y).int().numpy(), indexes.numpy(), cdf, ln, off)
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
        scales = self._scales_from(z_hat, mid_hw, y_hw)
        indexes = self.build_indexes(scales)
        cdf, off, ln = self._gc_tabs
        y_hat = torch.from_numpy(OR.decompress(strings[0], indexes.numpy(), cdf, ln, off)).float()
        s = self.g_s
        crop = lambda t, hw: t[..., : hw[0], : hw[1]]
        d5 = s.igdn5(crop(s.up5(y_hat), level_hw[3])); p5h = s.head5(d5)
        d4 = s.igdn4(crop(s.up4(d5), level_hw[2])); p4h = s.head4(d4)
        d3 = s.igdn3(crop(s.up3(d4), level_hw[1])); p3h = s.head3(d3)
        p2h = crop(s.up2(d3), level_hw[0])
        return {"x_hat": {"p2": p2h, "p3": p3h, "p4": p4h, "p5": p5h}}

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
        a = torch.relu(torch.randn(B, spec.F, h, w, generator=g)) * (1.0 + 0.5 * i)
        a = a + 0.1 * torch.randn(B, spec.F, h, w, generator=g)
        feats[k] = a.clamp_(-26.43, 28.40).to(device=device, dtype=dtype)
    return feats

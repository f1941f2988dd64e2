"""Generate tests/golden/*.npz from the REAL reference (CompressAI 1.2.4 vendored
in herok97/L-MSFC-V2-Training).

Runs only in the authoring container: it needs the reference built out-of-tree,

    cp -r /root/reference/compressai /tmp/refbuild/compressai
    chmod -R u+w /tmp/refbuild && cd /tmp/refbuild/compressai
    python setup.py build_ext --inplace          # 27 s

and is then invoked as  `python tests/golden/make_golden.py [/tmp/refbuild/compressai]`.
Nothing here is imported by tests or product code; the committed .npz files are
what travels.  Every array written is an output of reference code (modules from
compressai.layers / entropy_models / models.utils / losses, `_CXX`), evaluated
on seeded inputs that are stored alongside.
"""
import hashlib
import math
import os
import sys

import numpy as np
import torch
import torch.nn as nn

REF = sys.argv[1] if len(sys.argv) > 1 else "/tmp/refbuild/compressai"
sys.path.insert(0, REF)
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import compressai  # noqa: E402
from compressai._CXX import pmf_to_quantized_cdf as ref_cdf  # noqa: E402
from compressai.entropy_models import EntropyBottleneck, GaussianConditional  # noqa: E402
from compressai.layers import GDN  # noqa: E402
from compressai.losses import RateDistortionLoss  # noqa: E402
from compressai.models.base import get_scale_table  # noqa: E402
from compressai.models.utils import conv, deconv  # noqa: E402
from compressai.ops import LowerBound, quantize_ste  # noqa: E402
from compressai.optimizers import net_aux_optimizer  # noqa: E402

import importlib.util  # noqa: E402

_spec = importlib.util.spec_from_file_location(
    "lmsfc_spec", os.path.join(ROOT, "l-msfc-v2-training_b200", "spec.py"))
spec_mod = importlib.util.module_from_spec(_spec)
sys.modules["lmsfc_spec"] = spec_mod
_spec.loader.exec_module(spec_mod)
from oracle import lmsfc as OL  # noqa: E402

A = lambda t: t.detach().cpu().numpy().copy()


# ---------------------------------------------------------------------------
def gen_cdf():
    out = {}
    rng = np.random.default_rng(7)
    pmfs = [np.array([0.1, 0.2, 0.0, 0.0], np.float32),          # tests/test_ops.py:104-106
            np.array([0.5, 0.25, 0.25, 1e-9], np.float32)]
    for n in (2, 3, 5, 17, 64, 257, 1024):
        for kind in range(4):
            p = rng.random(n).astype(np.float32) ** (1 + 3 * kind)
            if kind >= 2:  # many exact zeros / tiny entries -> steal loop
                p[rng.random(n) < 0.5] = 0.0
                p[rng.integers(n)] = 1.0
            p = (p / p.sum()).astype(np.float32)
            pmfs.append(p)
    # un-normalised and very peaked rows
    pmfs.append(np.array([1e-7] * 30 + [1.0] + [1e-7] * 30, np.float32))
    pmfs.append((rng.random(100) * 3).astype(np.float32))
    flat = np.concatenate(pmfs)
    offs = np.cumsum([0] + [len(p) for p in pmfs]).astype(np.int64)
    cdfs = [np.asarray(ref_cdf(p.tolist(), 16), dtype=np.uint32) for p in pmfs]
    out["pmf_flat"], out["pmf_offsets"] = flat, offs
    out["cdf_flat"] = np.concatenate(cdfs)
    # precision 12 too
    out["cdf12_flat"] = np.concatenate(
        [np.asarray(ref_cdf(p.tolist(), 12), dtype=np.uint32) for p in pmfs[:10]])
    np.savez_compressed(os.path.join(HERE, "cdf.npz"), **out)
    print("cdf.npz", len(pmfs), "rows")


# ---------------------------------------------------------------------------
def gen_modules():
    out = {}
    torch.manual_seed(11)
    # LowerBound / STE (ops/bound_ops.py, ops/ops.py)
    x = torch.randn(64, requires_grad=True)
    g = torch.randn(64)
    lb = LowerBound(0.25)
    y = lb(x)
    y.backward(g)
    out["lb_x"], out["lb_g"], out["lb_y"], out["lb_gx"] = A(x), A(g), A(y), A(x.grad)
    x2 = (torch.randn(64) * 3).requires_grad_()
    y2 = quantize_ste(x2)
    y2.backward(g)
    out["ste_x"], out["ste_y"], out["ste_gx"] = A(x2), A(y2), A(x2.grad)

    # GDN / IGDN (layers/gdn.py:77-92) with perturbed parameters
    for name, inv in (("gdn", False), ("igdn", True)):
        C = 8
        m = GDN(C, inverse=inv)
        with torch.no_grad():
            m.beta.add_(torch.rand(C) * 0.5)
            m.gamma.add_(torch.rand(C, C) * 0.05)
            m.gamma[0, 1] = 0.0   # below the 2^-18 bound -> LowerBound gradient path
            m.beta[2] = 0.0
        x = (torch.randn(2, C, 5, 7) * 2).requires_grad_()
        gy = torch.randn(2, C, 5, 7)
        y = m(x)
        y.backward(gy)
        out[f"{name}_x"], out[f"{name}_gy"], out[f"{name}_y"] = A(x), A(gy), A(y)
        out[f"{name}_beta"], out[f"{name}_gamma"] = A(m.beta), A(m.gamma)
        out[f"{name}_gx"], out[f"{name}_gbeta"], out[f"{name}_ggamma"] = (
            A(x.grad), A(m.beta.grad), A(m.gamma.grad))
    # closed form at init, tests/test_layers.py:135-159
    m = GDN(4)
    x = torch.rand(1, 4, 3, 3)
    out["gdn_init_x"], out["gdn_init_y"] = A(x), A(m(x))

    # EntropyBottleneck (entropy_models.py:328-548)
    C = 6
    eb = EntropyBottleneck(C)
    with torch.no_grad():
        for n, p in eb.named_parameters():
            if n != "quantiles":
                p.add_(torch.randn_like(p) * 0.3)
        eb.quantiles.copy_(torch.tensor([-7.3, 0.4, 6.1]).repeat(C, 1, 1)
                           + torch.randn(C, 1, 3) * 0.5)
    for n, p in eb.named_parameters():
        out["eb_p_" + n] = A(p)
    out["eb_target"] = A(eb.target)
    x = (torch.randn(2, C, 4, 5) * 3).requires_grad_()
    eb.train()
    torch.manual_seed(123)
    xo, lik = eb(x)
    torch.manual_seed(123)
    noise = torch.empty(C, 1, 2 * 4 * 5).uniform_(-0.5, 0.5)  # entropy_models.py:161-165 draw
    glik = torch.randn_like(lik) * 0.1
    gout = torch.randn_like(xo)
    (torch.sum(torch.log(lik) * glik) + torch.sum(xo * gout)).backward()
    out["eb_x"], out["eb_noise"], out["eb_out"], out["eb_lik"] = A(x), A(noise), A(xo), A(lik)
    out["eb_glik"], out["eb_gout"], out["eb_gx"] = A(glik), A(gout), A(x.grad)
    for n, p in eb.named_parameters():
        out["eb_g_" + n] = A(p.grad) if p.grad is not None else np.zeros(0, np.float32)
    eb.zero_grad()
    eb.eval()
    xo_e, lik_e = eb(x.detach())
    out["eb_out_eval"], out["eb_lik_eval"] = A(xo_e), A(lik_e)
    aux = eb.loss()
    aux.backward()
    out["eb_aux"], out["eb_aux_gq"] = A(aux), A(eb.quantiles.grad)
    eb.update(force=True)
    out["eb_cdf"], out["eb_offset"], out["eb_cdf_length"] = (
        A(eb._quantized_cdf), A(eb._offset), A(eb._cdf_length))
    # a wide-input case: large |x| so the 1e-9 bound and its gradient rule fire
    xw = (torch.randn(1, C, 3, 3) * 40).requires_grad_()
    eb.train()
    torch.manual_seed(5)
    xo, lik = eb(xw)
    torch.manual_seed(5)
    noise = torch.empty(C, 1, 9).uniform_(-0.5, 0.5)
    torch.log(lik).sum().backward()
    out["ebw_x"], out["ebw_noise"], out["ebw_lik"], out["ebw_gx"] = A(xw), A(noise), A(lik), A(xw.grad)

    # GaussianConditional (entropy_models.py:551-689)
    gc = GaussianConditional(None)
    y = (torch.randn(2, 5, 4, 6) * 4).requires_grad_()
    sc = (torch.rand(2, 5, 4, 6) * 3).requires_grad_()
    with torch.no_grad():
        sc[0, 0] = 0.05          # below scale_bound 0.11
        y[0, 1] *= 20            # tails -> likelihood bound
    mu = torch.randn(2, 5, 4, 6).requires_grad_()
    gc.train()
    torch.manual_seed(77)
    yo, lik = gc(y, sc, mu)
    torch.manual_seed(77)
    noise = torch.empty_like(y).uniform_(-0.5, 0.5)
    glik = torch.randn_like(lik) * 0.1
    gout = torch.randn_like(yo)
    (torch.sum(torch.log(lik) * glik) + torch.sum(yo * gout)).backward()
    for k, v in dict(y=y, sc=sc, mu=mu, noise=noise, out=yo, lik=lik, glik=glik, gout=gout,
                     gy=y.grad, gsc=sc.grad, gmu=mu.grad).items():
        out["gc_" + k] = A(v)
    y.grad = sc.grad = None
    torch.manual_seed(78)
    yo, lik = gc(y, sc)          # no means
    torch.manual_seed(78)
    noise = torch.empty_like(y).uniform_(-0.5, 0.5)
    torch.log(lik).sum().backward()
    for k, v in dict(noise=noise, out=yo, lik=lik, gy=y.grad, gsc=sc.grad).items():
        out["gc0_" + k] = A(v)
    gc.eval()
    yo, lik = gc(y.detach(), sc.detach(), mu.detach())
    out["gc_out_eval"], out["gc_lik_eval"] = A(yo), A(lik)
    # table build: 8 scales stored fully, the default 64-scale table by hash + shape
    st8 = get_scale_table(0.11, 64, 8)
    gc.update_scale_table(st8, force=True)
    out["gc8_scale_table"], out["gc8_cdf"] = A(st8), A(gc._quantized_cdf)
    out["gc8_offset"], out["gc8_cdf_length"] = A(gc._offset), A(gc._cdf_length)
    st = get_scale_table()
    gc.update_scale_table(st, force=True)
    c = A(gc._quantized_cdf).astype(np.int32)
    out["gc64_scale_table"] = A(st)
    out["gc64_shape"] = np.array(c.shape)
    out["gc64_sha256"] = np.frombuffer(hashlib.sha256(c.tobytes()).digest(), dtype=np.uint8)
    out["gc64_offset"], out["gc64_cdf_length"] = A(gc._offset), A(gc._cdf_length)
    out["gc64_rows"] = c[[0, 13, 40, 63]]
    idx = gc.build_indexes(sc.detach())
    out["gc64_indexes"] = A(idx)

    # RateDistortionLoss (losses/rate_distortion.py:47-59)
    crit = RateDistortionLoss(lmbda=0.013)
    tgt = torch.rand(2, 3, 8, 8)
    xh = tgt + torch.randn_like(tgt) * 0.05
    l1, l2 = torch.rand(2, 4, 2, 2) * 0.9 + 0.05, torch.rand(2, 4, 1, 1) * 0.9 + 0.05
    r = crit({"x_hat": xh, "likelihoods": {"y": l1, "z": l2}}, tgt)
    out["rd_tgt"], out["rd_xh"], out["rd_l1"], out["rd_l2"] = A(tgt), A(xh), A(l1), A(l2)
    out["rd_out"] = np.array([r["loss"].item(), r["mse_loss"].item(), r["bpp_loss"].item()], np.float64)
    np.savez_compressed(os.path.join(HERE, "modules.npz"), **out)
    print("modules.npz", len(out), "arrays")


# ---------------------------------------------------------------------------
class RefLMSFC(nn.Module):
    """The frozen topology (spec.py) assembled from the REFERENCE's own modules."""

    def __init__(self, spec):
        super().__init__()
        self.spec = spec
        Fc, N, M = spec.F, spec.N, spec.M
        g_a = nn.Module()
        g_a.conv2, g_a.gdn2 = conv(Fc, N), GDN(N)
        g_a.conv3, g_a.gdn3 = conv(N + Fc, N), GDN(N)
        g_a.conv4, g_a.gdn4 = conv(N + Fc, N), GDN(N)
        g_a.conv5 = conv(N + Fc, M)
        self.g_a = g_a
        self.h_a = nn.Sequential(conv(M, N, stride=1, kernel_size=3), nn.ReLU(inplace=True),
                                 conv(N, N), nn.ReLU(inplace=True), conv(N, N))
        self.h_s = nn.Sequential(deconv(N, N), nn.ReLU(inplace=True), deconv(N, N),
                                 nn.ReLU(inplace=True), conv(N, M, stride=1, kernel_size=3),
                                 nn.ReLU(inplace=True))
        self.entropy_bottleneck = EntropyBottleneck(N)
        self.gaussian_conditional = GaussianConditional(None)
        g_s = nn.Module()
        g_s.up5, g_s.igdn5, g_s.head5 = deconv(M, N), GDN(N, inverse=True), conv(N, Fc, stride=1, kernel_size=3)
        g_s.up4, g_s.igdn4, g_s.head4 = deconv(N, N), GDN(N, inverse=True), conv(N, Fc, stride=1, kernel_size=3)
        g_s.up3, g_s.igdn3, g_s.head3 = deconv(N, N), GDN(N, inverse=True), conv(N, Fc, stride=1, kernel_size=3)
        g_s.up2 = deconv(N, Fc)
        self.g_s = g_s

    def forward(self, feats):
        crop = lambda t, r: t[..., : r.shape[-2], : r.shape[-1]]
        p2, p3, p4, p5 = (feats[k] for k in self.spec.levels)
        a, s = self.g_a, self.g_s
        e2 = a.gdn2(a.conv2(p2))
        e3 = a.gdn3(a.conv3(torch.cat([e2, p3], 1)))
        e4 = a.gdn4(a.conv4(torch.cat([e3, p4], 1)))
        y = a.conv5(torch.cat([e4, p5], 1))
        z = self.h_a(torch.abs(y))
        z_hat, z_lik = self.entropy_bottleneck(z)
        # h_s with every deconv output cropped to its target size (spec.py)
        hs = self.h_s
        z1 = self.h_a[2](self.h_a[1](self.h_a[0](torch.abs(y)))).shape
        t = hs[1](crop(hs[0](z_hat), torch.empty(z1)))
        t = hs[3](crop(hs[2](t), y))
        scales = hs[5](hs[4](t))
        y_hat, y_lik = self.gaussian_conditional(y, scales)
        d5 = s.igdn5(crop(s.up5(y_hat), p5)); p5h = s.head5(d5)
        d4 = s.igdn4(crop(s.up4(d5), p4)); p4h = s.head4(d4)
        d3 = s.igdn3(crop(s.up3(d4), p3)); p3h = s.head3(d3)
        p2h = crop(s.up2(d3), p2)
        return {"x_hat": {"p2": p2h, "p3": p3h, "p4": p4h, "p5": p5h},
                "likelihoods": {"y": y_lik, "z": z_lik},
                "latents": {"y": y, "z": z, "y_hat": y_hat, "z_hat": z_hat, "scales": scales}}


def ref_criterion(spec, out, feats, lmbda):
    """rate_distortion.py:47-59 applied level by level (adapter frozen in spec.py)."""
    lv = spec.levels
    B, _, H2, W2 = feats[lv[0]].shape
    num_pixels = spec.num_pixels(B, H2, W2)
    bpp = sum(torch.log(l).sum() / (-math.log(2) * num_pixels) for l in out["likelihoods"].values())
    mse_fn = nn.MSELoss()
    mse = sum(w * mse_fn(out["x_hat"][k], feats[k]) for w, k in zip(spec.level_weights, lv))
    return {"bpp_loss": bpp, "mse_loss": mse, "loss": lmbda * 255 ** 2 * mse + bpp}


def gen_step(tag, F, N, M, B, H2, W2, lmbda, seed):
    spec = spec_mod.LMSFCSpec(F=F, N=N, M=M)
    torch.manual_seed(seed)
    net = RefLMSFC(spec)
    with torch.no_grad():  # move off the init point so every gradient path is live
        for n, p in net.named_parameters():
            if "gdn" in n:
                p.add_(torch.rand_like(p) * 0.02)
            if "entropy_bottleneck._" in n:
                p.add_(torch.randn_like(p) * 0.1)
    net.train()
    feats = OL.synth_features(spec, B, H2, W2, seed=1234)
    out = {"cfg": np.array([F, N, M, B, H2, W2], np.int64), "lmbda": np.array(lmbda)}
    for k, v in feats.items():
        out["in_" + k] = A(v)
    sd0 = {k: v.clone() for k, v in net.state_dict().items()}
    for k, v in sd0.items():
        if v.numel() and v.dtype.is_floating_point and "reparam" not in k and "bound" not in k \
                and k.split(".")[-1] not in ("target",):
            out["w_" + k] = A(v)
    # optimisers exactly as examples/train.py:74-82
    conf = {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}
    opt = net_aux_optimizer(net, conf)
    opt["net"].zero_grad(); opt["aux"].zero_grad()
    # noise draws: EB first ([C,1,B*h*w]) then GC ([B,M,h,w]) -- google.py:295-297
    yshape = None
    torch.manual_seed(seed + 1)
    o = net(feats)
    zs, ys = o["latents"]["z"].shape, o["latents"]["y"].shape
    torch.manual_seed(seed + 1)
    nz = torch.empty(zs[1], 1, zs[0] * zs[2] * zs[3]).uniform_(-0.5, 0.5)
    ny = torch.empty(ys).uniform_(-0.5, 0.5)
    out["noise_z"], out["noise_y"] = A(nz), A(ny)
    crit = ref_criterion(spec, o, feats, lmbda)
    crit["loss"].backward()
    out["loss"] = np.array([crit["loss"].item(), crit["mse_loss"].item(), crit["bpp_loss"].item()], np.float64)
    for k in ("y", "z", "y_hat", "z_hat", "scales"):
        out["lat_" + k] = A(o["latents"][k])
    out["lik_y"], out["lik_z"] = A(o["likelihoods"]["y"]), A(o["likelihoods"]["z"])
    for k in spec.levels:
        out["xhat_" + k] = A(o["x_hat"][k])
    for n, p in net.named_parameters():
        if p.grad is not None:
            out["g_" + n] = A(p.grad)
    gn = torch.nn.utils.clip_grad_norm_(net.parameters(), 1.0)   # train.py:101-102
    out["grad_norm"] = A(gn)
    opt["net"].step()
    aux = net.entropy_bottleneck.loss()                            # train.py:105-107
    aux.backward()
    opt["aux"].step()
    out["aux_loss"] = A(aux)
    for n, p in net.named_parameters():
        out["w1_" + n] = A(p)
    np.savez_compressed(os.path.join(HERE, f"step_{tag}.npz"), **out)
    print(f"step_{tag}.npz loss/mse/bpp", out["loss"], "aux", float(aux), "gn", float(gn))


if __name__ == "__main__":
    print("reference:", compressai.__file__, "torch", torch.__version__)
    torch.set_num_threads(1)
    torch.use_deterministic_algorithms(True)
    gen_cdf()
    gen_modules()
    gen_step("tiny", 8, 8, 8, 2, 32, 32, 0.013, 42)
    gen_step("odd", 8, 16, 8, 1, 26, 38, 0.0067, 43)

"""GPU parity tests (run with -m gpu on a B200): every CUDA kernel, called through the C ABI via
the package's ctypes binding, against the oracle (oracle/ops.py: the restated reference maths,
autograd for gradients) evaluated in fp32 on the same device and the same seeded inputs.

Tolerances: tensor-core paths use bf16 operands with fp32 accumulation -> normalised error
(max|a-b| / max|b|) <= 2e-2 per tensor; fp32 entropy kernels rel 2e-4; integer CDFs bit-exact.
"""
import importlib
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def P():
    class NS:
        pass
    ns = NS()
    for m in ("functional", "layers", "entropy_models", "model", "losses", "optim", "train", "_lib", "spec"):
        setattr(ns, m, importlib.import_module(f"l-msfc-v2-training_b200.{m}"))
    return ns


def nerr(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


def bf(t):
    return t.to(torch.bfloat16).float()


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


# ---------------------------------------------------------------------------- conv / deconv
CONV_CASES = [
    # kind, k, s, c0, c1, cout, B, H, W, out_hw
    (0, 1, 1, 256, 0, 256, 1, 8, 16, None),       # plain GEMM, one tile
    (0, 3, 1, 256, 0, 256, 2, 9, 11, None),       # 3x3 head, ragged M
    (0, 5, 2, 256, 0, 256, 2, 32, 32, None),      # g_a.conv2-like
    (0, 5, 2, 256, 256, 256, 1, 13, 19, None),    # concat fusion, odd size (SURVEY §7 hard parts)
    (1, 5, 2, 256, 0, 256, 2, 8, 8, (16, 16)),    # deconv
    (1, 5, 2, 256, 0, 256, 1, 7, 10, (13, 19)),   # deconv cropped to an odd target
    (0, 5, 2, 256, 0, 256, 16, 64, 64, None),     # BLOCK_N = 256 path (M = 16384)
]


def _ref_conv(kind, k, s, x, w, b, out_hw):
    if kind == 0:
        return F.conv2d(x, w, b, stride=s, padding=k // 2)
    y = F.conv_transpose2d(x, w, b, stride=s, padding=k // 2, output_padding=s - 1)
    return y[..., : out_hw[0], : out_hw[1]]


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "k%d%ds%d_%d+%d-%d_b%d_%dx%d" % (c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7], c[8]))
def test_conv_fwd_bwd(P, case):
    kind, k, s, c0, c1, cout, B, H, W, out_hw = case
    cin = c0 + c1
    mod = (P.layers.Conv if kind == 0 else P.layers.Deconv)(cin, cout, k, s, **({"concat_channels": c1} if kind == 0 else {})).to(DEV)
    x0 = bf(rnd(B, c0, H, W, seed=1)).requires_grad_()
    x1 = bf(rnd(B, c1, H, W, seed=2)) if c1 else None
    with torch.no_grad():
        mod.weight.copy_(bf(mod.weight))
    # ---- product
    a0 = x0.detach().to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    y = mod(a0, x1, out_hw=out_hw)
    gy = bf(rnd(*y.shape, seed=3))
    y.backward(gy.to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
    torch.cuda.synchronize()
    # ---- oracle (fp32 torch on the same device)
    w = mod.weight.detach().clone().contiguous().requires_grad_()
    b = mod.bias.detach().clone().requires_grad_()
    xin = torch.cat([x0, x1], 1) if c1 else x0
    yr = _ref_conv(kind, k, s, xin, w, b, out_hw)
    yr.backward(gy)
    assert y.shape == yr.shape
    e = {"y": nerr(y, yr), "gx": nerr(a0.grad, x0.grad), "gw": nerr(mod.weight.grad, w.grad), "gb": nerr(mod.bias.grad, b.grad)}
    print(case, e)
    assert all(v < 2e-2 for v in e.values()), e


def test_conv_relu_and_latent_head(P):
    mod = P.layers.Conv(256, 256, 3, 1).to(DEV)
    x = bf(rnd(2, 256, 6, 5, seed=5))
    y = mod(x, relu=True)
    yr = F.relu(F.conv2d(x, bf(mod.weight.detach()), mod.bias, padding=1))
    assert nerr(y, yr) < 2e-2
    yf, ya = mod(x, latent_head=True)
    yr = F.conv2d(x, bf(mod.weight.detach()), mod.bias, padding=1)
    assert yf.dtype == torch.float32 and nerr(yf, yr) < 1e-2 and nerr(ya, yr.abs()) < 2e-2
    # gradient through both outputs: g_direct + sign(y) * g_abs
    xa = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    yf, ya = mod(xa, latent_head=True)
    g1, g2 = bf(rnd(*yf.shape, seed=6)), bf(rnd(*yf.shape, seed=7))
    torch.autograd.backward([yf, ya], [g1, g2.to(torch.bfloat16)])
    xr = x.clone().requires_grad_()
    yr = F.conv2d(xr, bf(mod.weight.detach()), mod.bias, padding=1)
    torch.autograd.backward([yr, yr.abs()], [g1, g2])
    assert nerr(xa.grad, xr.grad) < 2e-2


def test_conv_leaky_relu_and_plain_latent_head(P):
    """The mean-scale hyperprior's pieces (models/google.py:377-398): LeakyReLU(0.01) in the conv epilogue and its
    gradient from the output's sign; a latent head that returns (fp32 y, bf16 y) whose two gradients add."""
    mod = P.layers.Deconv(256, 256).to(DEV)
    x = bf(rnd(2, 256, 5, 4, seed=8))
    xa = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    y = mod(xa, relu="leaky")
    g = bf(rnd(*y.shape, seed=9))
    y.backward(g.to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
    xr = x.clone().requires_grad_()
    w = bf(mod.weight.detach()).requires_grad_()
    yr = F.leaky_relu(F.conv_transpose2d(xr, w, mod.bias, stride=2, padding=2, output_padding=1), 0.01)
    yr.backward(g)
    assert float((y.float() < 0).float().mean()) > 0.2          # the negative branch is exercised
    assert nerr(y, yr) < 2e-2 and nerr(xa.grad, xr.grad) < 2e-2 and nerr(mod.weight.grad, w.grad) < 2e-2
    head = P.layers.Conv(256, 512, 3, 1).to(DEV)
    xa = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    yf, yb = head(xa, latent_head="plain")
    yr = F.conv2d(x, bf(head.weight.detach()), head.bias, padding=1)
    assert yf.dtype == torch.float32 and yb.dtype == torch.bfloat16 and nerr(yf, yr) < 1e-2 and nerr(yb, yr) < 2e-2
    g1, g2 = bf(rnd(*yf.shape, seed=6)), bf(rnd(*yf.shape, seed=7))
    torch.autograd.backward([yf, yb], [g1, g2.to(torch.bfloat16)])
    xr = x.clone().requires_grad_()
    yr = F.conv2d(xr, bf(head.weight.detach()), head.bias, padding=1)
    yr.backward(g1 + g2)
    assert nerr(xa.grad, xr.grad) < 2e-2


# ---------------------------------------------------------------------------- GDN
@pytest.mark.parametrize("inverse", [False, True])
@pytest.mark.parametrize("shape", [(2, 256, 5, 7), (4, 256, 32, 32)])
def test_gdn_fwd_bwd(P, inverse, shape):
    from oracle import ops as O
    mod = P.layers.GDN(shape[1], inverse=inverse).to(DEV)
    with torch.no_grad():   # move gamma off its init so the off-diagonal terms matter
        mod.gamma.add_(0.02 * torch.rand_like(mod.gamma))
        mod.beta.add_(0.1 * torch.rand_like(mod.beta))
    x = bf(rnd(*shape, seed=11, scale=2.0))
    gy = bf(rnd(*shape, seed=12))
    xa = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    y = mod(xa)
    y.backward(gy.to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
    xr = x.clone().requires_grad_()
    beta = mod.beta.detach().clone().requires_grad_()
    gamma = mod.gamma.detach().clone().requires_grad_()
    yr = O.gdn(xr, beta, gamma, inverse=inverse)
    yr.backward(gy)
    e = {"y": nerr(y, yr), "gx": nerr(xa.grad, xr.grad), "gbeta": nerr(mod.beta.grad, beta.grad),
         "ggamma": nerr(mod.gamma.grad, gamma.grad)}
    print(inverse, shape, e)
    assert all(v < 3e-2 for v in e.values()), e


def test_gdn_closed_form_at_init(P):
    # compressai/tests/test_layers.py:135-159: y == x / sqrt(1 + 0.1 x^2) at init (IGDN: x * sqrt)
    x = bf(rnd(1, 256, 3, 3, seed=13))
    y = P.layers.GDN(256).to(DEV)(x)
    assert nerr(y, x / torch.sqrt(1 + 0.1 * x ** 2)) < 1e-2
    y = P.layers.GDN(256, inverse=True).to(DEV)(x)
    assert nerr(y, x * torch.sqrt(1 + 0.1 * x ** 2)) < 1e-2


# ---------------------------------------------------------------------------- entropy models
def test_entropy_bottleneck_golden(P, golden):
    g = golden("modules")
    C = g["eb_x"].shape[1]
    eb = P.entropy_models.EntropyBottleneck(C).to(DEV)
    with torch.no_grad():
        for n, p in eb.named_parameters():
            p.copy_(torch.from_numpy(g["eb_p_" + n]))
    x = torch.from_numpy(g["eb_x"]).to(DEV).requires_grad_()
    B, _, H, W = x.shape
    noise = torch.from_numpy(g["eb_noise"]).to(DEV).reshape(C, B, H, W).permute(1, 0, 2, 3).contiguous()
    out, lik = eb(x, training=True, noise=noise)
    T = lambda k: torch.from_numpy(g[k]).to(DEV)
    (torch.sum(torch.log(lik) * T("eb_glik")) + torch.sum(out * T("eb_gout"))).backward()
    np.testing.assert_allclose(out.detach().cpu().numpy(), g["eb_out"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(lik.detach().cpu().numpy(), g["eb_lik"], rtol=2e-4, atol=1e-7)
    np.testing.assert_allclose(x.grad.cpu().numpy(), g["eb_gx"], rtol=2e-3, atol=2e-5)
    for n, p in eb.named_parameters():
        if n == "quantiles":
            continue
        np.testing.assert_allclose(p.grad.cpu().numpy(), g["eb_g_" + n], rtol=2e-3, atol=2e-5, err_msg=n)
    # aux loss + gradient wrt quantiles (entropy_models.py:431-434)
    eb.zero_grad()
    aux = eb.loss()
    aux.backward()
    np.testing.assert_allclose(float(aux), float(g["eb_aux"]), rtol=1e-5)
    np.testing.assert_allclose(eb.quantiles.grad.cpu().numpy(), g["eb_aux_gq"], rtol=1e-4, atol=1e-6)
    # eval mode: dequantize around the medians
    with torch.no_grad():
        out, lik = eb(x.detach(), training=False)
    np.testing.assert_allclose(out.cpu().numpy(), g["eb_out_eval"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(lik.cpu().numpy(), g["eb_lik_eval"], rtol=2e-4, atol=1e-7)
    # integer tables against the reference's own (bit-exact)
    assert eb.update(force=True)
    assert np.array_equal(eb._quantized_cdf.cpu().numpy(), g["eb_cdf"])
    assert np.array_equal(eb._offset.cpu().numpy(), g["eb_offset"])
    assert np.array_equal(eb._cdf_length.cpu().numpy(), g["eb_cdf_length"])


def test_gaussian_conditional_golden(P, golden):
    import hashlib
    g = golden("modules")
    T = lambda k: torch.from_numpy(g[k]).to(DEV)
    gc = P.entropy_models.GaussianConditional(None).to(DEV)
    y, sc, mu = (T(k).requires_grad_() for k in ("gc_y", "gc_sc", "gc_mu"))
    out, lik = gc(y, sc, mu, training=True, noise=T("gc_noise"))
    (torch.sum(torch.log(lik) * T("gc_glik")) + torch.sum(out * T("gc_gout"))).backward()
    np.testing.assert_allclose(out.detach().cpu().numpy(), g["gc_out"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(lik.detach().cpu().numpy(), g["gc_lik"], rtol=2e-4, atol=1e-9)
    np.testing.assert_allclose(y.grad.cpu().numpy(), g["gc_gy"], rtol=2e-3, atol=2e-5)
    np.testing.assert_allclose(sc.grad.cpu().numpy(), g["gc_gsc"], rtol=2e-3, atol=2e-5)
    np.testing.assert_allclose(mu.grad.cpu().numpy(), g["gc_gmu"], rtol=2e-3, atol=2e-5)
    out, lik = gc(y.detach(), sc.detach(), mu.detach(), training=False)
    np.testing.assert_allclose(out.cpu().numpy(), g["gc_out_eval"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(lik.cpu().numpy(), g["gc_lik_eval"], rtol=2e-4, atol=1e-9)
    # CDF tables: 8-scale table exact, 64-scale table by hash + sampled rows (as tests/test_oracle_golden.py)
    gc.update_scale_table([float(v) for v in g["gc8_scale_table"]], force=True)
    assert np.array_equal(gc._quantized_cdf.cpu().numpy(), g["gc8_cdf"])
    assert np.array_equal(gc._offset.cpu().numpy(), g["gc8_offset"])
    gc.update_scale_table(torch.from_numpy(g["gc64_scale_table"]), force=True)
    cdf = gc._quantized_cdf.cpu().numpy().astype(np.int32)
    assert list(cdf.shape) == g["gc64_shape"].tolist()
    assert hashlib.sha256(cdf.tobytes()).digest() == g["gc64_sha256"].tobytes()
    assert np.array_equal(cdf[[0, 13, 40, 63]], g["gc64_rows"])
    assert np.array_equal(gc._cdf_length.cpu().numpy(), g["gc64_cdf_length"])


@pytest.mark.parametrize("with_means", [False, True])
def test_gaussian_conditional_vs_oracle(P, with_means):
    from oracle import ops as O
    shape = (3, 16, 7, 9)
    y = rnd(*shape, seed=21, scale=3.0).requires_grad_()
    s = (rnd(*shape, seed=22).abs() * 2 + 0.01).requires_grad_()     # some scales below the 0.11 bound
    mu = rnd(*shape, seed=23).requires_grad_() if with_means else None
    noise = torch.rand(shape, generator=torch.Generator().manual_seed(24)).to(DEV) - 0.5
    gl, go = rnd(*shape, seed=25), rnd(*shape, seed=26)
    gc = P.entropy_models.GaussianConditional(None).to(DEV)
    out, lik = gc(y, s, mu, training=True, noise=noise)
    torch.autograd.backward([out, lik], [go, gl])
    got = [t.grad.clone() for t in (y, s) + ((mu,) if with_means else ())]
    for t in (y, s) + ((mu,) if with_means else ()):
        t.grad = None
    outr, likr = O.gc_forward(y, s, mu, True, noise)
    torch.autograd.backward([outr, likr], [go, gl])
    ref = [t.grad for t in (y, s) + ((mu,) if with_means else ())]
    torch.testing.assert_close(out, outr, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(lik, likr, rtol=2e-4, atol=1e-7)
    for a, b in zip(got, ref):
        torch.testing.assert_close(a, b, rtol=2e-3, atol=2e-5)
    # fused sum(ln L)
    assert abs(float(getattr(lik, P.functional._SUM_LN_ATTR)) - float(torch.log(likr).sum())) < 1e-3 * abs(float(torch.log(likr).sum()))
    # eval mode
    out, lik = gc(y.detach(), s.detach(), None if mu is None else mu.detach(), training=False)
    outr, likr = O.gc_forward(y.detach(), s.detach(), None if mu is None else mu.detach(), False)
    torch.testing.assert_close(out, outr, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(lik, likr, rtol=2e-4, atol=1e-7)


def test_philox_noise_statistics(P):
    gc = P.entropy_models.GaussianConditional(None).to(DEV)
    y = torch.zeros(4, 64, 32, 32, device=DEV)
    s = torch.ones_like(y)
    o1, _ = gc(y, s, training=True)
    o2, _ = gc(y, s, training=True)
    assert float(o1.min()) >= -0.5 and float(o1.max()) < 0.5
    assert abs(float(o1.mean())) < 5e-3 and abs(float(o1.var()) - 1 / 12) < 5e-3
    assert not torch.equal(o1, o2)          # the device-side step counter advanced
    with pytest.raises(ValueError):
        P.functional.gc_apply(y, s, None, None, "bogus", 0.11, 1e-9)


# ---------------------------------------------------------------------------- loss / optimiser / CDF
def test_rd_loss_vs_oracle(P):
    from oracle import ops as O
    spec = P.spec.DEFAULT_SPEC
    shapes = [(2, 256, 8, 8), (2, 256, 4, 4), (2, 256, 2, 2), (2, 256, 1, 1)]
    xh = [bf(rnd(*s, seed=30 + i)).requires_grad_() for i, s in enumerate(shapes)]
    xt = [bf(rnd(*s, seed=40 + i)) for i, s in enumerate(shapes)]
    liks = [torch.rand(2, 256, 1, 1, device=DEV).clamp_min(1e-3).requires_grad_() for _ in range(2)]
    lmbda, npix = 0.013, spec.num_pixels(2, 8, 8)
    xa = [t.detach().to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_() for t in xh]
    la = [t.detach().clone().requires_grad_() for t in liks]
    loss, bpp, mse = P.functional.rd_loss_apply(xa, xt, la, lmbda * 255 ** 2, npix, spec.level_weights)
    loss.backward()
    ref = O.rd_loss(liks, xh, xt, lmbda, npix, spec.level_weights)
    ref["loss"].backward()
    assert abs(float(loss) - float(ref["loss"])) < 1e-3 * abs(float(ref["loss"]))
    assert abs(float(bpp) - float(ref["bpp_loss"])) < 1e-4 * abs(float(ref["bpp_loss"]))
    assert abs(float(mse) - float(ref["mse_loss"])) < 1e-3 * abs(float(ref["mse_loss"]))
    for a, b in zip(xa, xh):
        assert nerr(a.grad, b.grad) < 1e-2
    for a, b in zip(la, liks):
        torch.testing.assert_close(a.grad, b.grad, rtol=1e-5, atol=1e-9)


def test_rd_loss_single_launch_vs_oracle(P):
    """lmsfc_rd_loss: the likelihood kernels' sum(ln L) scalars, every level's squared error and gradient, and the final
    bpp / D / loss (+ a scalar riding along) in one launch; twice in a row gives bit-identical numbers (fixed-order
    reduction, nothing to zero beforehand)."""
    from oracle import ops as O
    spec = P.spec.DEFAULT_SPEC
    shapes = [(3, 256, 40, 24), (3, 256, 20, 12), (3, 256, 10, 6), (3, 256, 5, 3)]
    xh = [bf(rnd(*s, seed=50 + i)).requires_grad_() for i, s in enumerate(shapes)]
    xt = [bf(rnd(*s, seed=60 + i)) for i, s in enumerate(shapes)]
    y = rnd(3, 256, 3, 2, seed=70, scale=3.0)
    scales = torch.rand(3, 256, 3, 2, device=DEV) * 2 + 0.05
    noise = torch.rand(3, 256, 3, 2, device=DEV) - 0.5
    lmbda, npix = 0.013, spec.num_pixels(3, 40, 24)
    extra = torch.tensor(123.5, device=DEV)
    vals = []
    for _ in range(2):
        xa = [t.detach().to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_() for t in xh]
        ya = y.clone().requires_grad_()
        _, lik = P.functional.gc_apply(ya, scales, None, noise, "noise", 0.11, 1e-9)
        loss, bpp, mse, s4 = P.functional.rd_loss_apply(xa, xt, [lik], lmbda * 255 ** 2, npix, spec.level_weights,
                                                         extra=extra, with_scalars=True)
        assert s4 is not None
        loss.backward()
        torch.cuda.synchronize()
        vals.append([float(v) for v in s4] + [float(ya.grad.abs().sum())])
    assert vals[0] == vals[1]
    assert vals[0][:3] == [float(loss), float(bpp), float(mse)] and vals[0][3] == 123.5
    yr = y.clone().requires_grad_()
    _, lik_r = O.gc_forward(yr, scales, None, True, noise)
    ref = O.rd_loss([lik_r], xh, xt, lmbda, npix, spec.level_weights)
    ref["loss"].backward()
    assert abs(float(loss) - float(ref["loss"])) < 1e-3 * abs(float(ref["loss"]))
    assert abs(float(bpp) - float(ref["bpp_loss"])) < 1e-4 * abs(float(ref["bpp_loss"]))
    assert abs(float(mse) - float(ref["mse_loss"])) < 1e-3 * abs(float(ref["mse_loss"]))
    for a, b in zip(xa, xh):
        assert nerr(a.grad, b.grad) < 1e-2
    assert nerr(ya.grad, yr.grad) < 1e-3


# 10003: the register-staged kernel with a scalar tail; 148 * 2048 * 3 + 1237: the bulk-copy pipeline (adam_stream_kernel,
# one persistent CTA per SM, whole 2048-parameter chunks) with a ragged remainder handled by its last CTA
@pytest.mark.parametrize("n", [10000 + 3, 148 * 2048 * 3 + 1237])
def test_fused_clip_adam_vs_torch(P, n):
    p0, g0 = rnd(n, seed=50), rnd(n, seed=51, scale=0.05)
    p = p0.clone()
    g = g0.clone()
    opt = P.optim.FusedAdam(p, g, lr=1e-2, clip_max_norm=1.0, bf16_shadow=torch.empty(n, dtype=torch.bfloat16, device=DEV))
    pr = torch.nn.Parameter(p0.clone())
    ropt = torch.optim.Adam([pr], lr=1e-2)
    for it in range(3):
        g.copy_(g0 * (it + 1))
        pr.grad = (g0 * (it + 1)).clone()
        torch.nn.utils.clip_grad_norm_([pr], 1.0)
        ropt.step()
        opt.step()
    torch.testing.assert_close(p, pr.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(opt.bf16_shadow.float(), p.to(torch.bfloat16).float())
    assert abs(float(opt.grad_norm) - float((g0 * 3).norm())) < 1e-3 * float((g0 * 3).norm())
    # the same update slice by slice with the gradient cleared in the same pass (the pipelined step's form)
    g.copy_(g0)
    pr.grad = g0.clone()
    torch.nn.utils.clip_grad_norm_([pr], 1.0)
    ropt.step()
    opt.accumulate_norm()
    opt.prep()
    cut = (n // 3) // 64 * 64
    opt.update(0, cut, zero_grad=True)
    opt.update(cut, n, zero_grad=True)
    torch.testing.assert_close(p, pr.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(opt.bf16_shadow.float(), p.to(torch.bfloat16).float())
    assert float(g.abs().max()) == 0.0


def test_cdf_device_bit_exact(P, golden):
    g = golden("cdf")
    offs = g["pmf_offsets"]
    rows = len(offs) - 1
    lens = np.diff(offs) - 1                       # last entry of every row is its tail mass
    maxlen = int(lens.max())
    pmf = np.zeros((rows, maxlen), np.float32)
    tail = np.zeros(rows, np.float32)
    want = np.zeros((rows, maxlen + 2), np.int32)
    pos = 0
    for i in range(rows):
        r = g["pmf_flat"][offs[i]:offs[i + 1]]
        pmf[i, : lens[i]], tail[i] = r[:-1], r[-1]
        want[i, : len(r) + 1] = g["cdf_flat"][pos:pos + len(r) + 1]
        pos += len(r) + 1
    L = P._lib
    d = lambda a: torch.from_numpy(a).to(DEV)
    pmf_d, tail_d, len_d = d(pmf), d(tail), d(lens.astype(np.int32))
    cdf = torch.empty((rows, maxlen + 2), dtype=torch.int32, device=DEV)
    st = torch.zeros(rows, dtype=torch.int32, device=DEV)
    L.pmf_rows_to_cdf(L.ptr(pmf_d), maxlen, L.ptr(tail_d), L.ptr(len_d), rows, 16, L.ptr(cdf), maxlen + 2, L.ptr(st),
                      L.stream())
    assert int(st.max()) == 0
    assert np.array_equal(cdf.cpu().numpy(), want)


def test_update_tables_match_oracle(P):
    from oracle import cdf as OC
    eb = P.entropy_models.EntropyBottleneck(32).to(DEV)
    assert eb.update() and not eb.update()
    m, b, f = ([getattr(eb, f"_{k}{i}").detach().cpu() for i in range(n)] for k, n in (("matrix", 5), ("bias", 5), ("factor", 4)))
    cdf, off, ln = OC.eb_update(m, b, f, eb.quantiles.detach().cpu())
    assert np.array_equal(eb._quantized_cdf.cpu().numpy(), cdf)
    assert np.array_equal(eb._offset.cpu().numpy(), off) and np.array_equal(eb._cdf_length.cpu().numpy(), ln)
    gc = P.entropy_models.GaussianConditional(None).to(DEV)
    table = P.model.get_scale_table()
    assert gc.update_scale_table(table)
    cdf, off, ln = OC.gc_update(table)
    assert np.array_equal(gc._quantized_cdf.cpu().numpy(), cdf)
    assert np.array_equal(gc._offset.cpu().numpy(), off) and np.array_equal(gc._cdf_length.cpu().numpy(), ln)


def test_device_pmf_grids_and_tables(P):
    """update() with the float pmf evaluated on the device (eb_pmf / gc_pmf kernels) against the default
    host evaluation (bit-identical to the reference).  The fp32 grids agree to a few ulp (erfc / tanh / exp of libdevice
    vs the CPU), so the rounded 16-bit counts agree except where pmf * 2^16 sits on a rounding boundary (the Gaussian
    table has entries at EXACTLY x.5).  One flipped count changes the row total, hence the rescale and the order of the
    zero-width "steal" loop (ops.cpp:74-100): the CDF positions of that row can then move by many counts -- which is why
    the host evaluation stays the default (tables bit-identical to CompressAI's) and the device one is opt-in for
    per-frame refresh, where encoder and decoder build their tables the same way.  Checked here: lengths / offsets
    equal, count-level agreement of the grids, every device-built row a valid CDF, coder round trip on device tables."""
    import time
    from oracle import cdf as OC
    L = P._lib
    torch.manual_seed(3)
    eb_h = P.entropy_models.EntropyBottleneck(256).to(DEV)
    with torch.no_grad():                       # "trained" parameters: move everything off its init
        for n, p in eb_h.named_parameters():
            if n == "quantiles":
                p.copy_(torch.tensor([-12.0, 0.3, 9.0]).repeat(256, 1, 1).to(DEV) + torch.randn(256, 1, 3, device=DEV) * 2)
                p[:, 0, 0].clamp_(max=-1.0); p[:, 0, 2].clamp_(min=1.5); p[:, 0, 1].clamp_(-0.9, 1.4)
            else:
                p.add_(0.3 * torch.randn_like(p))
    eb_d = P.entropy_models.EntropyBottleneck(256).to(DEV)
    eb_d.load_state_dict(eb_h.state_dict())
    eb_d.device_pmf = True
    res = {}
    for name, m in (("host", eb_h), ("device", eb_d)):
        m.update(force=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            m.update(force=True)
        torch.cuda.synchronize()
        res["eb_ms_" + name] = round((time.perf_counter() - t0) / 5 * 1e3, 3)
    assert torch.equal(eb_h._offset, eb_d._offset) and torch.equal(eb_h._cdf_length, eb_d._cdf_length)
    d = (eb_h._quantized_cdf - eb_d._quantized_cdf).abs()
    rows_off = int((d.max(1).values > 0).sum())
    res["eb_rows_differing"] = (rows_off, int(d.shape[0]), int(d.max()))
    assert rows_off <= 0.1 * d.shape[0]
    gc_h = P.entropy_models.GaussianConditional(None).to(DEV)
    gc_d = P.entropy_models.GaussianConditional(None).to(DEV)
    gc_d.device_pmf = True
    table = P.model.get_scale_table()
    for name, m in (("host", gc_h), ("device", gc_d)):
        m.update_scale_table(table, force=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            m.update_scale_table(table, force=True)
        torch.cuda.synchronize()
        res["gc_ms_" + name] = round((time.perf_counter() - t0) / 5 * 1e3, 3)
    assert torch.equal(gc_h._offset, gc_d._offset) and torch.equal(gc_h._cdf_length, gc_d._cdf_length)
    d = (gc_h._quantized_cdf - gc_d._quantized_cdf).abs()
    res["gc_rows_differing"] = (int((d.max(1).values > 0).sum()), int(d.shape[0]), int(d.max()))
    # the float grid itself, against the oracle's restatement of entropy_models.py:625-648
    pmf_o, tail_o, center, plen, ml = OC.gc_pmf(np.asarray(table, dtype=np.float32))
    plen = plen.numpy()
    pmf_dv = torch.empty((len(plen), ml), device=DEV)
    tail_dv = torch.empty(len(plen), device=DEV)
    L.gc_pmf(L.ptr(gc_d.scale_table), L.ptr(center.int().to(DEV)), len(plen), ml, L.ptr(pmf_dv), ml, L.ptr(tail_dv), L.stream())
    a, b = pmf_o.numpy().astype(np.float32), pmf_dv.cpu().numpy()
    ca, cb = np.floor(a * np.float32(65536) + np.float32(0.5)), np.floor(b * np.float32(65536) + np.float32(0.5))
    valid = np.arange(ml)[None, :] < np.asarray(plen)[:, None]
    flips = int(((ca != cb) & valid).sum())
    res["gc_count_flips"] = (flips, int(valid.sum()), float(np.abs(ca - cb)[valid].max()))
    assert flips <= 1e-3 * valid.sum() and np.abs(ca - cb)[valid].max() <= 1
    assert float(np.abs(a - b)[valid].max()) < 5e-7          # a few ulp of the erfc values (~0.5) whose difference this is
    print("update() ms, rows whose CDF differs (rows, of, max count distance), Gaussian count flips:", res)
    # a table is valid whichever way it was built: strictly increasing rows ending at 2^16, and the coder round-trips on it
    for m in (eb_d, gc_d):
        cdf, ln = m._quantized_cdf, m._cdf_length
        for r in range(cdf.shape[0]):
            row = cdf[r, : int(ln[r])]
            assert int(row[0]) == 0 and int(row[-1]) == 1 << 16 and bool((row[1:] > row[:-1]).all())
    x = torch.round(torch.randn(2, 256, 4, 4, device=DEV) * 3)
    strings = eb_d.compress(x)
    back = eb_d.decompress(strings, x.shape[-2:])
    med = eb_d._get_medians().detach().reshape(1, -1, 1, 1)
    assert torch.equal(back, torch.round(x - med) + med)

"""The GDN / IGDN contraction fused behind the GEMM that produces its input (csrc/igemm.cu, FUSE kernels), through the
C ABI: lmsfc_conv_gdn_forward against lmsfc_conv_forward -> lmsfc_gdn_forward of the same library (same bf16 roundings:
x and x^2 must be bit-identical) and against the oracle in fp32; lmsfc_conv_dgrad_gdn_backward against
lmsfc_conv_dgrad (+ addend) -> lmsfc_gdn_backward and against autograd through the oracle's GDN (layers/gdn.py:77-92).
"""
import importlib

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"
PKG = "l-msfc-v2-training_b200"
BF16, CL = torch.bfloat16, torch.channels_last


@pytest.fixture(scope="module")
def L():
    return importlib.import_module(f"{PKG}._lib")


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def nhwc(t):
    return t.to(BF16).contiguous(memory_format=CL)


def nerr(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-20))


def _gdn_params(C, seed):
    g = torch.Generator().manual_seed(seed)
    gamma = (0.1 * torch.eye(C) + 0.02 * torch.rand(C, C, generator=g)).to(DEV)
    beta = (1.0 + 0.1 * torch.rand(C, generator=g)).to(DEV)
    return gamma.to(BF16).contiguous(), beta.contiguous()


def _weights(kind, k, cin, cout, seed):
    """bf16 GEMM-layout weight [D0][T][D1] and its torch-layout fp32 view."""
    d0, d1 = (cout, cin) if kind == 0 else (cin, cout)
    w = rnd(d0, k * k, d1, seed=seed, scale=(cin * k * k) ** -0.5).to(BF16)
    wt = w.float().view(d0, k, k, d1).permute(0, 3, 1, 2).contiguous()
    return w, wt


def _ref_conv(kind, k, s, x, w, b, out_hw):
    if kind == 0:
        return F.conv2d(x, w, b, stride=s, padding=k // 2)
    y = F.conv_transpose2d(x, w, b, stride=s, padding=k // 2, output_padding=s - 1)
    return y[..., : out_hw[0], : out_hw[1]]


FWD_CASES = [
    # kind, k, s, c0, c1, B, H, W, out_hw, inverse
    (0, 5, 2, 256, 0, 16, 64, 64, None, False),        # g_a.conv2 + gdn2 at the bench size
    (1, 5, 2, 256, 0, 16, 16, 16, (32, 32), True),     # g_s.up3 + igdn3 at the bench size (4 sub-pixel phases)
    (0, 5, 2, 256, 256, 1, 13, 19, None, False),       # concat fusion, ragged tile
    (1, 5, 2, 256, 0, 1, 7, 10, (13, 19), True),       # cropped deconv
    (0, 5, 2, 256, 256, 1, 100, 152, None, False),     # cfg-4 conv3 + gdn3
]


@pytest.mark.parametrize("case", FWD_CASES, ids=lambda c: "k%d_%d+%d_b%d_%dx%d_inv%d" % (c[0], c[3], c[4], c[5], c[6], c[7], c[9]))
def test_conv_gdn_forward_fused(L, case):
    from oracle import ops as O
    kind, k, s, c0, c1, B, H, W, out_hw, inverse = case
    C = 256
    Ho, Wo = ((H + s - 1) // s, (W + s - 1) // s) if kind == 0 else out_hw
    d = L.ConvDesc(kind, k, s, B, H, W, Ho, Wo, c0, c1, C)
    x0 = nhwc(rnd(B, c0, H, W, seed=1))
    x1 = nhwc(rnd(B, c1, H, W, seed=2)) if c1 else None
    w, wt = _weights(kind, k, c0 + c1, C, seed=3)
    bias = rnd(C, seed=4, scale=0.1)
    gamma, beta = _gdn_params(C, 5)
    P, st = L.ptr, L.stream()
    mk = lambda: torch.empty((B, C, Ho, Wo), dtype=BF16, device=DEV, memory_format=CL)
    # ---- unfused: conv -> (square) -> GDN GEMM
    xu, x2u, yu, rnu = mk(), mk(), mk(), mk()
    L.conv_forward(d, P(x0), P(x1), P(w), P(bias), 0, P(xu), None, None, st)
    L.gdn_forward(P(xu), P(gamma), P(beta), int(inverse), B * Ho * Wo, C, P(x2u), P(yu), P(rnu), st)
    # ---- fused
    xf, x2f, yf, rnf = (torch.full_like(xu, float("nan")) for _ in range(4))
    L.conv_gdn_forward(d, P(x0), P(x1), P(w), P(bias), P(gamma), P(beta), int(inverse), P(xf), P(x2f), P(yf), P(rnf), st)
    torch.cuda.synchronize()
    if L.gdn_fusable(d, inference=True):              # the separate conv uses the same whole-K 256-wide tiles: identical summation order
        assert torch.equal(xf, xu), nerr(xf, xu)
        assert torch.equal(x2f, x2u), nerr(x2f, x2u)
    else:                             # it splits K over a cluster: last-bit differences of the bf16 rounding
        assert nerr(xf, xu) < 1e-2 and nerr(x2f, x2u) < 2e-2
    assert torch.equal(x2f.float(), (xf.float() ** 2).to(BF16).float())
    e = {"y_vs_unfused": nerr(yf, yu), "rn_vs_unfused": nerr(rnf, rnu)}
    # ---- oracle, fp32
    xin = torch.cat([x0.float(), x1.float()], 1) if c1 else x0.float()
    xr = _ref_conv(kind, k, s, xin, wt, bias, out_hw)
    g32, b32 = gamma.float(), beta
    nrm = F.conv2d(xr ** 2, g32.view(C, C, 1, 1), b32)
    yr = xr * (torch.sqrt(nrm) if inverse else torch.rsqrt(nrm))
    e.update(x=nerr(xf, xr), y=nerr(yf, yr), rn=nerr(rnf, torch.rsqrt(nrm)))
    print(case, e)
    assert e["y_vs_unfused"] < 1e-2 and e["rn_vs_unfused"] < 1e-2, e
    assert e["x"] < 2e-2 and e["y"] < 2e-2 and e["rn"] < 2e-2, e
    # inference form: only y is written
    y2 = torch.full_like(yu, float("nan"))
    L.conv_gdn_forward(d, P(x0), P(x1), P(w), P(bias), P(gamma), P(beta), int(inverse), None, None, P(y2), None, st)
    torch.cuda.synchronize()
    assert torch.equal(y2, yf)


BWD_CASES = [
    # the conv whose data gradient produces grad(y): kind, k, s, c0, c1, cout, B, H, W (its INPUT = the GDN's output), out_hw, inverse, addend
    (0, 5, 2, 256, 256, 256, 16, 32, 32, None, False, False),   # g_a.conv3's dgrad + gdn2's backward at the bench size
    (1, 5, 2, 256, 0, 256, 16, 32, 32, (64, 64), True, True),   # g_s.up2's dgrad (+ head3's) + igdn3's backward
    (0, 3, 1, 256, 0, 256, 2, 9, 11, None, True, False),        # a 3x3 head's dgrad, ragged
    (1, 5, 2, 256, 0, 256, 1, 13, 19, (26, 38), True, True),    # odd sizes
    (0, 5, 2, 256, 256, 256, 1, 25, 38, None, False, False),    # cfg-4 conv5's dgrad + gdn4
]


@pytest.mark.parametrize("case", BWD_CASES, ids=lambda c: "k%d%d_b%d_%dx%d_inv%d_add%d" % (c[0], c[1], c[6], c[7], c[8], c[10], c[11]))
def test_conv_dgrad_gdn_backward_fused(L, case):
    from oracle import ops as O
    kind, k, s, c0, c1, cout, B, H, W, out_hw, inverse, with_add = case
    C = 256
    Ho, Wo = ((H + s - 1) // s, (W + s - 1) // s) if kind == 0 else out_hw
    d = L.ConvDesc(kind, k, s, B, H, W, Ho, Wo, c0, c1, cout)
    Pn = B * H * W
    w, wt = _weights(kind, k, c0 + c1, cout, seed=13)
    gout = nhwc(rnd(B, cout, Ho, Wo, seed=14))
    add = nhwc(rnd(B, C, H, W, seed=15, scale=0.5)) if with_add else None
    gamma, beta = _gdn_params(C, 16)
    # the GDN's forward state: x (its input) and rn, from the library's own forward
    x = nhwc(rnd(B, C, H, W, seed=17, scale=2.0))
    P, st = L.ptr, L.stream()
    mk = lambda: torch.empty((B, C, H, W), dtype=BF16, device=DEV, memory_format=CL)
    x2, y, rn = mk(), mk(), mk()
    L.gdn_forward(P(x), P(gamma), P(beta), int(inverse), Pn, C, P(x2), P(y), P(rn), st)
    # ---- unfused: dgrad -> (+ addend) -> gdn_t -> GEMM
    gy = mk()
    L.conv_dgrad(d, P(gout), P(w), C, P(gy), st)
    if with_add:
        gy = (gy.float() + add.float()).to(BF16).contiguous(memory_format=CL)
    tu, gxu = mk(), mk()
    L.gdn_backward(P(gy), P(x), P(rn), P(gamma), int(inverse), Pn, C, P(x2), 1, P(tu), P(gxu), None, None, st)
    # ---- fused
    tf_, gxf = torch.full_like(tu, float("nan")), torch.full_like(tu, float("nan"))
    L.conv_dgrad_gdn_backward(d, P(gout), P(w), C, P(add), P(x), P(rn), P(gamma), int(inverse), P(tf_), P(gxf), st)
    torch.cuda.synchronize()
    e = {"t_vs_unfused": nerr(tf_, tu), "gx_vs_unfused": nerr(gxf, gxu)}
    # ---- oracle: autograd through conv(GDN(x)) in fp32
    xr = x.float().clone().requires_grad_()
    g32 = gamma.float()
    nrm = F.conv2d(xr ** 2, g32.view(C, C, 1, 1), beta)
    yr = xr * (torch.sqrt(nrm) if inverse else torch.rsqrt(nrm))
    if kind == 0:
        src = torch.cat([yr, torch.zeros(B, c1, H, W, device=DEV)], 1) if c1 else yr
    else:
        src = yr
    out = _ref_conv(kind, k, s, src, wt, None, out_hw)
    loss = (out * gout.float()).sum()
    if with_add:
        loss = loss + (yr * add.float()).sum()
    loss.backward()
    e["gx"] = nerr(gxf, xr.grad)
    print(case, e)
    assert e["t_vs_unfused"] < 2e-2 and e["gx_vs_unfused"] < 2e-2, e
    assert e["gx"] < 3e-2, e


def test_fusable_follows_the_tiling_cost_model(L):
    big = L.ConvDesc(0, 5, 2, 16, 64, 64, 32, 32, 256, 0, 256)
    small = L.ConvDesc(0, 5, 2, 16, 16, 16, 8, 8, 256, 256, 256)
    assert L.gdn_fusable(big, inference=True) and not L.gdn_fusable(small, inference=True)
    assert L.gdn_fusable(L.ConvDesc(0, 5, 2, 16, 32, 32, 16, 16, 256, 256, 256), dgrad=True, n_in=256)
    wide = L.ConvDesc(0, 3, 1, 16, 32, 32, 32, 32, 256, 0, 512)
    assert not L.gdn_fusable(wide, inference=True)

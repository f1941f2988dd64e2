"""Graph-timed A/B of the fused conv+GDN kernels against the separate launches (bench shapes, solo, warm).
  python profiles/fused_bench.py"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
L = importlib.import_module("l-msfc-v2-training_b200._lib")
dev = torch.device("cuda", 0)
BF16, CL = torch.bfloat16, torch.channels_last
P = L.ptr


def timeit(fn, n=10, reps=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / n * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def t(*shape):
    return torch.randn(*shape, device=dev).to(BF16).contiguous(memory_format=CL)


C = 256
gamma = (0.1 * torch.eye(C) + 0.02 * torch.rand(C, C)).to(dev).to(BF16).contiguous()
beta = torch.ones(C, device=dev)
bias = torch.zeros(C, device=dev)
for name, (kind, k, s, c0, c1, B, H, W, Ho, Wo, inv) in {
        "conv2+gdn2": (0, 5, 2, 256, 0, 16, 64, 64, 32, 32, 0), "up3+igdn3": (1, 5, 2, 256, 0, 16, 16, 16, 32, 32, 1),
        "conv3+gdn3": (0, 5, 2, 256, 256, 16, 32, 32, 16, 16, 0), "up4+igdn4": (1, 5, 2, 256, 0, 16, 8, 8, 16, 16, 1)}.items():
    d = L.ConvDesc(kind, k, s, B, H, W, Ho, Wo, c0, c1, C)
    x0, x1 = t(B, c0, H, W), (t(B, c1, H, W) if c1 else None)
    d0, d1 = (C, c0 + c1) if kind == 0 else (c0 + c1, C)
    w = (torch.randn(d0, k * k, d1, device=dev) * 0.02).to(BF16)
    x, x2, y, rn = (torch.empty((B, C, Ho, Wo), dtype=BF16, device=dev, memory_format=CL) for _ in range(4))
    st = L.stream
    tc = timeit(lambda: L.conv_forward(d, P(x0), P(x1), P(w), P(bias), 0, P(x), None, None, st()))
    tu = timeit(lambda: (L.conv_forward(d, P(x0), P(x1), P(w), P(bias), 0, P(x), None, None, st()),
                         L.gdn_forward(P(x), P(gamma), P(beta), inv, B * Ho * Wo, C, P(x2), P(y), P(rn), st())))
    tf = timeit(lambda: L.conv_gdn_forward(d, P(x0), P(x1), P(w), P(bias), P(gamma), P(beta), inv, P(x), P(x2), P(y), P(rn), st()))
    ti = timeit(lambda: L.conv_gdn_forward(d, P(x0), P(x1), P(w), P(bias), P(gamma), P(beta), inv, None, None, P(y), None, st()))
    print(f"fwd {name:12s} conv {tc:6.1f}  conv->gdn {tu:6.1f}  fused {tf:6.1f}  fused(y only) {ti:6.1f} us  fusable={L.gdn_fusable(d)}")
for name, (kind, k, s, c0, c1, co, B, H, W, Ho, Wo, inv, add) in {
        "conv3.dgrad+gdn2": (0, 5, 2, 256, 256, 256, 16, 32, 32, 16, 16, 0, 0), "up2.dgrad+igdn3": (1, 5, 2, 256, 0, 256, 16, 32, 32, 64, 64, 1, 1),
        "conv4.dgrad+gdn3": (0, 5, 2, 256, 256, 256, 16, 16, 16, 8, 8, 0, 0), "up3.dgrad+igdn4": (1, 5, 2, 256, 0, 256, 16, 16, 16, 32, 32, 1, 1)}.items():
    d = L.ConvDesc(kind, k, s, B, H, W, Ho, Wo, c0, c1, co)
    d0, d1 = (co, c0 + c1) if kind == 0 else (c0 + c1, co)
    w = (torch.randn(d0, k * k, d1, device=dev) * 0.02).to(BF16)
    gout = t(B, co, Ho, Wo)
    x, rn, x2, gy, tt, gx, addt = (t(B, C, H, W) for _ in range(7))
    rn.abs_().add_(0.1)
    st = L.stream
    Pn = B * H * W
    tc = timeit(lambda: L.conv_dgrad(d, P(gout), P(w), C, P(gy), st()))
    tu = timeit(lambda: (L.conv_dgrad(d, P(gout), P(w), C, P(gy), st()),
                         L.gdn_backward(P(gy), P(x), P(rn), P(gamma), inv, Pn, C, P(x2), 1, P(tt), P(gx), None, None, st())))
    tf = timeit(lambda: L.conv_dgrad_gdn_backward(d, P(gout), P(w), C, P(addt) if add else None, P(x), P(rn), P(gamma), inv, P(tt), P(gx), st()))
    print(f"bwd {name:17s} dgrad {tc:6.1f}  dgrad->gdn_bwd {tu:6.1f}  fused {tf:6.1f} us  fusable={L.gdn_fusable(d, True, 256)}")

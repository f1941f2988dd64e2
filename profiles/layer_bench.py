"""Per-layer timing of the L-MSFC GEMM kernels through the C ABI (CUDA events, warm, averaged) at the
bench workload (batch 16, P2 64x64): forward, data-gradient and weight-gradient of every conv /
deconv of spec.py plus GDN forward / backward.  Prints one table; `--json` writes it for profiles/.

  python profiles/layer_bench.py [--batch 16] [--h2 64] [--w2 64] [--reps 30]
"""
import argparse
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--h2", type=int, default=64)
    ap.add_argument("--w2", type=int, default=64)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    L = importlib.import_module(PKG + "._lib")
    layers = importlib.import_module(PKG + ".layers")
    Fn = importlib.import_module(PKG + ".functional")
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def timeit(fn, cold=False, n=10):
        """fn captured n times back to back into one CUDA graph (host/ctypes/descriptor-encode time
        is then off the clock, as in the captured train step); median over replays, per call."""
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n):
                fn()
        ts = []
        for _ in range(a.reps):
            if cold:
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / n)
        ts.sort()
        return ts[len(ts) // 2]

    B = a.batch
    rows = []
    rs = torch.zeros(2, dtype=torch.int64, device=dev)
    print(f"per-launch floor (1-thread kernel, in a graph): {timeit(lambda: L.rng_advance(L.ptr(rs), L.stream())):.2f} us")
    bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    P = L.ptr
    for (name, kind, k, ci, co, hi, wi, ho, wo) in spec.conv_layers(a.h2, a.w2):
        if kind == "gdn":
            inv = int("igdn" in name)
            x, gy = bf(B, ci, hi, wi), bf(B, ci, hi, wi)
            x2, y, rn, t, gx = (torch.empty_like(x) for _ in range(5))
            gam = (torch.eye(ci, device=dev) * 0.1 + 1e-3).to(torch.bfloat16)
            beta = torch.ones(ci, device=dev)
            gg, gbt = torch.zeros(ci, ci, device=dev), torch.zeros(ci, device=dev)
            Pn = B * hi * wi
            fw = lambda: L.gdn_forward(P(x), P(gam), P(beta), inv, Pn, ci, P(x2), P(y), P(rn), L.stream())
            bw = lambda: L.gdn_backward(P(gy), P(x), P(rn), P(gam), inv, Pn, ci, P(x2), 1, P(t), P(gx), P(gg), P(gbt), L.stream())
            fw()
            t_f, t_fc, t_b = timeit(fw), timeit(fw, cold=True, n=1), timeit(bw)
            fl = 2.0 * ci * co * hi * wi * B
            rows.append((name, f"{B}x{hi}x{wi}", fl, t_f, t_fc, t_b, None))
            continue
        s = 1 if k == 3 else 2
        c1 = 256 if (kind == "conv" and ci == 512) else 0
        m = (layers.Conv if kind == "conv" else layers.Deconv)(ci, co, k, s, concat_channels=c1).to(dev)
        st = m._state
        x0 = bf(B, ci - c1, hi, wi)
        x1 = bf(B, c1, hi, wi) if c1 else None
        st.refresh(m.weight)
        d = st.desc(B, hi, wi, ho, wo)
        yb = Fn.empty_nhwc(B, co, ho, wo, torch.bfloat16, dev)
        gy = bf(B, co, ho, wo)
        gx = Fn.empty_nhwc(B, ci - c1, hi, wi, torch.bfloat16, dev)
        gw = torch.zeros((st.D0, st.T, st.D1), device=dev)
        gb = torch.zeros(co, device=dev)
        wf = wd = st.w_direct
        f = lambda: L.conv_forward(d, P(x0), P(x1), P(wf), P(m.bias), 0, P(yb), None, None, L.stream())
        dg = lambda: L.conv_dgrad(d, P(gy), P(wd), ci - c1, P(gx), L.stream())
        wg = lambda: L.conv_wgrad(d, P(x0), P(x1), P(gy), P(gw), P(gb), L.stream())
        px = ho * wo if kind == "conv" else hi * wi
        fl = 2.0 * k * k * ci * co * px * B
        rows.append((name, f"{B}x{hi}x{wi}->{ho}x{wo} k{k} {ci}->{co}", fl, timeit(f), timeit(f, cold=True, n=1), timeit(dg), timeit(wg)))
    print(f"{'layer':12s} {'shape':34s} {'GFLOP':>7s} {'fwd us':>8s} {'(cold)':>8s} {'dgrad/bwd':>9s} {'wgrad':>8s} {'fwd TF/s':>9s}")
    tot = [0.0, 0.0, 0.0, 0.0]
    for (n, sh, fl, tf, tfc, tb, tw) in rows:
        print(f"{n:12s} {sh:34s} {fl/1e9:7.2f} {tf:8.1f} {tfc:8.1f} {tb:9.1f} {(tw if tw is not None else 0):8.1f} {fl/tf/1e6:9.1f}")
        tot[0] += fl; tot[1] += tf; tot[2] += tb; tot[3] += tw or 0
    print(f"{'total':12s} {'':34s} {tot[0]/1e9:7.2f} {tot[1]:8.1f} {'':8s} {tot[2]:9.1f} {tot[3]:8.1f}   sum {tot[1]+tot[2]+tot[3]:.1f} us; "
          f"3x flops at 1396 TF/s = {3*tot[0]/1396e6:.1f} us")
    if a.json:
        json.dump([dict(layer=n, shape=sh, flops=fl, fwd_us=tf, fwd_cold_us=tfc, bwd_us=tb, wgrad_us=tw) for (n, sh, fl, tf, tfc, tb, tw) in rows],
                  open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()

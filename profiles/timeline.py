"""%globaltimer timeline of the implicit-GEMM kernel (debug build: make -C .../csrc TIMELINE=1).
  LMSFC_LIB=l-msfc-v2-training_b200/libb200lmsfc_tl.so python profiles/timeline.py
Prints, for a few layer shapes, per-CTA stage times (us since the first CTA started)."""
import ctypes, importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("LMSFC_LIB", os.path.join(ROOT, "l-msfc-v2-training_b200", "libb200lmsfc_tl.so"))
PKG = "l-msfc-v2-training_b200"
L = importlib.import_module(PKG + "._lib")
layers = importlib.import_module(PKG + ".layers")
Fn = importlib.import_module(PKG + ".functional")
dev = torch.device("cuda", 0)
NAMES = ["start", "setup", "gather-issued", "gather-landed", "tmem-full", "reds-issued", "fence+bar", "epi-done", "exit", "mma-first", "mma-last"]
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
P = L.ptr
CASES = [("h_a.4 16x2x2->1x1 k5", 0, 5, 2, 256, 16, 2, 2, 1, 1), ("h_a.0 16x4x4 k3", 0, 3, 1, 256, 16, 4, 4, 4, 4),
         ("head5 16x8x8 k3", 0, 3, 1, 256, 16, 8, 8, 8, 8), ("conv4 16x16x16->8x8 k5 (256 in)", 0, 5, 2, 256, 16, 16, 16, 8, 8),
         ("conv2 16x64x64->32x32 k5", 0, 5, 2, 256, 16, 64, 64, 32, 32)]
for (name, kind, k, s, c, B, hi, wi, ho, wo) in CASES:
    m = layers.Conv(c, 256, k, s).to(dev)
    st = m._state
    st.refresh(m.weight)
    x, y = bf(B, c, hi, wi), Fn.empty_nhwc(B, 256, ho, wo, torch.bfloat16, dev)
    d = st.desc(B, hi, wi, ho, wo)
    for _ in range(3):
        L.conv_forward(d, P(x), None, P(st.w_direct), P(m.bias), 0, P(y), None, None, L.stream())
    torch.cuda.synchronize()
    buf = np.zeros((64, 16), dtype=np.uint64)
    L.lib.lmsfc_debug_read_timeline(ctypes.c_void_p(buf.ctypes.data))
    t0 = buf[:, 0][buf[:, 0] > 0].min()
    print(f"== {name}")
    for cta in (0, 1, 31, 63):
        if buf[cta, 0] == 0:
            continue
        r = buf[cta]
        print(f"  cta {cta:2d}: " + "  ".join(f"{NAMES[i]} {((int(r[i]) - int(t0)) / 1e3):.2f}" for i in (0, 1, 9, 2, 3, 10, 4, 5, 6, 7, 8) if r[i] >= t0))
    print(f"  last exit over recorded CTAs: {(int(buf[:, 8].max()) - int(t0)) / 1e3:.2f} us;  SM clock over cta 0's MMA loop "
          f"{(int(buf[0, 12]) - int(buf[0, 11])) / max(1, int(buf[0, 10]) - int(buf[0, 9])) * 1e3:.0f} MHz")

# ---- weight-gradient kernel (same debug build): stage stamps of a few CTAs
WNAMES = {0: "start", 1: "setup", 9: "mma-first", 10: "mma-last-issued", 4: "tmem-full", 5: "reds-issued", 7: "reds-done", 8: "exit"}
for (name, k, s, B, hi, wi, ho, wo) in [("conv2 wgrad", 5, 2, 16, 64, 64, 32, 32), ("head3 wgrad", 3, 1, 16, 32, 32, 32, 32),
                                        ("head5 wgrad", 3, 1, 16, 8, 8, 8, 8)]:
    m = layers.Conv(256, 256, k, s).to(dev)
    st = m._state
    st.refresh(m.weight)
    d = st.desc(B, hi, wi, ho, wo)
    x, gy = bf(B, 256, hi, wi), bf(B, 256, ho, wo)
    gw = torch.zeros((st.D0, st.T, st.D1), device=dev)
    for _ in range(3):
        L.conv_wgrad(d, P(x), None, P(gy), P(gw), None, L.stream())
    torch.cuda.synchronize()
    buf = np.zeros((64, 16), dtype=np.uint64)
    L.lib.lmsfc_debug_read_wg_timeline(ctypes.c_void_p(buf.ctypes.data))
    t0 = buf[:, 0][buf[:, 0] > 0].min()
    print(f"== {name}")
    for cta in (0, 1, 31, 63):
        r = buf[cta]
        if r[0] == 0:
            continue
        print(f"  cta {cta:2d}: " + "  ".join(f"{WNAMES[i]} {((int(r[i]) - int(t0)) / 1e3):.2f}" for i in (0, 1, 9, 10, 4, 5, 7, 8) if r[i] >= t0))
    print(f"  last exit over recorded CTAs: {(int(buf[:, 8].max()) - int(t0)) / 1e3:.2f} us;  SM clock over cta 0's MMA loop "
          f"{(int(buf[0, 12]) - int(buf[0, 11])) / max(1, int(buf[0, 10]) - int(buf[0, 9])) * 1e3:.0f} MHz")

# ---- GDN forward / backward GEMMs (same implicit-GEMM kernel, EPI_GDN_FWD / EPI_GDN_BWD)
for (name, B, h, w) in [("gdn 16x32x32", 16, 32, 32), ("gdn 16x8x8", 16, 8, 8)]:
    C, Pn = 256, B * h * w
    x, gy = bf(B, C, h, w), bf(B, C, h, w)
    gam = (torch.rand(C, C, device=dev) * 0.01).to(torch.bfloat16)
    beta = torch.ones(C, device=dev)
    x2, y, rn, t, gx = (torch.empty_like(x) for _ in range(5))
    gg, gb = torch.zeros(C, C, device=dev), torch.zeros(C, device=dev)
    for which in ("fwd", "bwd"):
        for _ in range(3):
            if which == "fwd":
                L.gdn_forward(P(x), P(gam), P(beta), 0, Pn, C, P(x2), P(y), P(rn), L.stream())
            else:
                L.gdn_backward(P(gy), P(x), P(rn), P(gam), 0, Pn, C, P(x2), 1, P(t), P(gx), None, P(gb), L.stream())
        torch.cuda.synchronize()
        buf = np.zeros((64, 16), dtype=np.uint64)
        L.lib.lmsfc_debug_read_timeline(ctypes.c_void_p(buf.ctypes.data))
        t0 = buf[:, 0][buf[:, 0] > 0].min()
        print(f"== {name} {which}")
        EN = dict(enumerate(NAMES)); EN.update({13: "epi:inputs-staged", 14: "epi:slab0-computed", 15: "epi:slab0-stored", 3: "epi:bwd-issued", 5: "epi:bwd-landed", 2: "epi:tmem0", 6: "epi:tmem1"})
        for cta in (0, 7):
            r = buf[cta]
            if r[0] == 0:
                continue
            print(f"  cta {cta:2d}: " + "  ".join(f"{EN[i]} {((int(r[i]) - int(t0)) / 1e3):.2f}" for i in (0, 1, 9, 10, 4, 13, 3, 5, 2, 6, 14, 15, 7, 8) if r[i] >= t0))

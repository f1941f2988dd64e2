"""Timing experiments on the weight-gradient kernel (debug build, make TIMELINE=1): LMSFC_WG_DBG bit0 skips the
gather, bit1 skips the TMA loads -> which producer bounds the main loop?  Usage: LMSFC_WG_DBG=n python profiles/wgrad_probe.py"""
import importlib, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("LMSFC_LIB", os.path.join(ROOT, "l-msfc-v2-training_b200", "libb200lmsfc_tl.so"))
PKG = "l-msfc-v2-training_b200"
L = importlib.import_module(PKG + "._lib"); layers = importlib.import_module(PKG + ".layers")
dev = torch.device("cuda", 0)
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
P = L.ptr
for (name, k, s, B, hi, wi, ho, wo) in [("conv2", 5, 2, 16, 64, 64, 32, 32), ("head3", 3, 1, 16, 32, 32, 32, 32)]:
    m = layers.Conv(256, 256, k, s).to(dev); st = m._state; st.refresh(m.weight)
    d = st.desc(B, hi, wi, ho, wo)
    x, gy = bf(B, 256, hi, wi), bf(B, 256, ho, wo)
    gw = torch.zeros((st.D0, st.T, st.D1), device=dev)
    y = torch.empty_like(gy); gx = torch.empty_like(x)
    def t(fn, n=10):
        side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side): fn()
        torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n): fn()
        ts = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3 / n)
        ts.sort(); return ts[len(ts)//2]
    tw = t(lambda: L.conv_wgrad(d, P(x), None, P(gy), P(gw), None, L.stream()))
    tf = t(lambda: L.conv_forward(d, P(x), None, P(st.w_direct), P(m.bias), 0, P(y), None, None, L.stream()))
    td = t(lambda: L.conv_dgrad(d, P(gy), P(st.w_direct), 256, P(gx), L.stream()))
    print(f"LMSFC_WG_DBG={os.environ.get('LMSFC_WG_DBG','0')} {name}: wgrad {tw:.1f} us  fwd {tf:.1f} us  dgrad {td:.1f} us (graph-timed, per launch)")
    import ctypes, numpy as np
    def loop(fn_read):        # cta 0: MMA loop duration in ns and SM cycles (stamps of the last launch)
        buf = np.zeros((64, 16), dtype=np.uint64); fn_read(ctypes.c_void_p(buf.ctypes.data)); r = buf[0]
        return int(r[10]) - int(r[9]), int(r[12]) - int(r[11])
    L.conv_wgrad(d, P(x), None, P(gy), P(gw), None, L.stream()); torch.cuda.synchronize()
    ns, cyc = loop(L.lib.lmsfc_debug_read_wg_timeline)
    print(f"   wgrad cta0 MMA loop after sustained run: {ns/1e3:.1f} us, {cyc} cyc ({cyc/ns*1e3:.0f} MHz)")
    L.conv_forward(d, P(x), None, P(st.w_direct), P(m.bias), 0, P(y), None, None, L.stream()); torch.cuda.synchronize()
    ns, cyc = loop(L.lib.lmsfc_debug_read_timeline)
    print(f"   fwd   cta0 MMA loop after sustained run: {ns/1e3:.1f} us, {cyc} cyc ({cyc/ns*1e3:.0f} MHz)")

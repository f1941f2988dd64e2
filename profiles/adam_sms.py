"""How many SMs does the optimiser pass need?  adam_stream_kernel (csrc/optim.cu: one persistent bulk-copy CTA per SM) on N
SMs, N from LMSFC_ADAM_STREAM (0: the plain adam_kernel); one child process per N (the setting is read once).  Prints GB/s
over 27.9 M parameters x 34 B and checks one update against the same formula in torch.
    python profiles/adam_sms.py [N ...]"""
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"


def child():
    import torch
    L = importlib.import_module(PKG + "._lib")
    P, dev = L.ptr, "cuda"
    n = 27_900_000 // 64 * 64
    p_, m_, v_, g_ = (torch.zeros(n, dtype=torch.float32, device=dev) for _ in range(4))
    g_.normal_(0, 1e-3)
    sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
    scratch = torch.zeros(8, dtype=torch.float32, device=dev)
    step = torch.zeros(1, dtype=torch.int32, device=dev)
    L.adam_prep(1e-4, None, 0.9, 0.999, 0.0, None, P(step), P(scratch), None, 1, None, L.stream())
    f = lambda: L.adam_update(P(p_), P(m_), P(v_), P(g_), P(sh), n, 0.9, 0.999, 1e-8, P(scratch), 1, L.stream())
    # one update from a known state against the same formula in torch (fp32)
    m_.normal_(0, 1e-3); v_.uniform_(0, 1e-6); p_.normal_(0, 0.1)
    g0, m0, v0, p0 = g_.clone(), m_.clone(), v_.clone(), p_.clone()
    f()
    torch.cuda.synchronize()
    sc = scratch.cpu()
    clip, bc1, bc2, lr = float(sc[0]), float(sc[1]), float(sc[2]), float(sc[4])
    gg = g0 * clip
    m1 = 0.9 * m0 + (1 - 0.9) * gg
    v1 = 0.999 * v0 + (1 - 0.999) * gg * gg
    p1 = p0 - (lr / bc1) * m1 / (v1.sqrt() * (bc2 ** -0.5) + 1e-8)
    err = max(float((m_ - m1).abs().max() / m1.abs().max()), float((v_ - v1).abs().max() / v1.abs().max()),
              float((p_ - p1).abs().max() / p1.abs().max()))
    ok = err < 1e-5 and float(g_.abs().max()) == 0.0 and torch.equal(sh, p_.to(torch.bfloat16))
    g_.copy_(g0)
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    ms = ts[len(ts) // 2]
    sms = os.environ.get("LMSFC_ADAM_STREAM", "148")
    kind = "plain" if sms == "0" else "stream"
    sms = "all" if sms == "0" else sms
    print(f"adam {kind:>6} on {sms:>4} SMs: {ms * 1e3:8.1f} us  {n * 34 / ms / 1e6:8.1f} GB/s  "
          f"({n * 34 / ms / 1e6 / (int(sms) if sms != 'all' else 148):6.1f} GB/s per SM)  check {'ok' if ok else 'FAILED'} "
          f"(err {err:.1e})", flush=True)


if __name__ == "__main__":
    if os.environ.get("ADAM_SMS_CHILD"):
        child()
    else:
        ns = sys.argv[1:] or ["0", "8", "16", "20", "24", "32", "48", "64", "96", "148"]
        for n in ns:
            env = dict(os.environ, ADAM_SMS_CHILD="1", LMSFC_ADAM_STREAM=n)
            subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)

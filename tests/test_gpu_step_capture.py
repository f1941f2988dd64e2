"""lmsfc_step_capture_begin / _end / _launch / _destroy (include/b200lmsfc.h): a chain of C-ABI calls recorded once and
replayed on rewritten inputs must produce exactly what the same calls produce eagerly (same kernels, same grids ->
bit-identical), incl. a fork onto a second stream inside the capture.  No torch graph object is involved: the caller
owns every buffer, the library owns the graph."""
import ctypes
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF16, F32 = torch.bfloat16, torch.float32
CL = torch.channels_last


@pytest.fixture(scope="module")
def M():
    class NS:
        pass
    ns = NS()
    for m in ("_lib", "layers", "functional"):
        setattr(ns, m, importlib.import_module(f"l-msfc-v2-training_b200.{m}"))
    return ns


def test_captured_chain_replays_like_eager(M):
    L, P = M._lib, M._lib.ptr
    torch.manual_seed(3)
    B, H, W, C = 2, 32, 32, 256
    conv = M.layers.Conv(C, C, 5, 2).to(DEV)
    gdn = M.layers.GDN(C).to(DEV)
    x = torch.randn(B, C, H, W, device=DEV).to(BF16).contiguous(memory_format=CL)
    with torch.no_grad():                              # builds the bf16 GEMM-layout weight and warms every kernel up
        conv(x)
    st = conv._state
    d = st.desc(B, H, W, H // 2, W // 2)
    Pn = B * (H // 2) * (W // 2)
    beta_eff = torch.empty(C, dtype=F32, device=DEV)
    gamma_bf = torch.empty(C, C, dtype=BF16, device=DEV)
    y = torch.empty(B, C, H // 2, W // 2, device=DEV, dtype=BF16).contiguous(memory_format=CL)
    x2, out, rn, grad = (torch.empty_like(y) for _ in range(4))
    target = torch.randn_like(y)
    sse = torch.zeros(1, dtype=F32, device=DEV)
    colsum = torch.zeros(C, dtype=F32, device=DEV)
    side = torch.cuda.Stream()
    main = torch.cuda.Stream()

    def chain(s_main, s_side):
        """reparam -> conv -> GDN -> [MSE + gradient on the main stream || column sums of the GDN output on a forked one]"""
        L.reparam_forward(P(gdn.beta), C, gdn._state.beta_bound, M.functional.PEDESTAL, P(beta_eff), None, s_main.cuda_stream)
        L.reparam_forward(P(gdn.gamma), C * C, gdn._state.gamma_bound, M.functional.PEDESTAL, None, P(gamma_bf), s_main.cuda_stream)
        L.conv_forward(d, P(x), None, P(st.w_direct), P(conv.bias), 0, P(y), None, None, s_main.cuda_stream)
        L.gdn_forward(P(y), P(gamma_bf), P(beta_eff), 0, Pn, C, P(x2), P(out), P(rn), s_main.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(s_main)
        s_side.wait_event(ev)
        L.colsum_static(P(out), Pn, C, P(colsum), s_side.cuda_stream)
        L.mse_forward_backward(P(out), P(target), out.numel(), 0.5, P(sse), P(grad), s_main.cuda_stream)
        ev2 = torch.cuda.Event()
        ev2.record(s_side)
        s_main.wait_event(ev2)

    def reset():
        sse.zero_(); colsum.zero_()
        for t in (y, x2, out, rn, grad):
            t.zero_()

    # eager results for two different inputs
    inputs = [x.clone(), (torch.randn_like(x.float()) * 2).to(BF16).contiguous(memory_format=CL)]
    want = []
    main.wait_stream(torch.cuda.current_stream())
    for xin in inputs:
        x.copy_(xin)
        reset()
        torch.cuda.synchronize()
        chain(main, side)
        torch.cuda.synchronize()
        want.append([t.clone() for t in (out, grad, sse, colsum, rn)])

    reset()
    torch.cuda.synchronize()
    handle, nk = ctypes.c_void_p(), ctypes.c_int(0)
    before = L.launch_count()
    L.step_capture_begin(main.cuda_stream)
    chain(main, side)
    L.step_capture_end(main.cuda_stream, ctypes.byref(handle), ctypes.byref(nk))
    n_k = L.launch_count() - before                     # reparam x2, conv, GDN (square + GEMM), column sums, MSE
    assert handle.value and nk.value == n_k and n_k >= 6, (nk.value, n_k)
    torch.cuda.synchronize()
    assert float(sse) == 0.0 and float(out.float().abs().max()) == 0.0, "capture must record, not execute"
    for xin, w in zip(inputs + inputs[:1], want + want[:1]):
        x.copy_(xin)
        reset()
        torch.cuda.synchronize()
        L.step_launch(handle, main.cuda_stream)
        torch.cuda.synchronize()
        for name, got, ref in zip(("out", "grad", "sse", "colsum", "rn"), (out, grad, sse, colsum, rn), w):
            if name == "sse":                          # one float atomic per block: the order of the adds is not fixed
                assert abs(float(got) - float(ref)) <= 1e-5 * abs(float(ref)), (float(got), float(ref))
            else:
                assert torch.equal(got, ref), name
    assert L.launch_count() - before == n_k + 3 * n_k
    L.step_destroy(handle)
    L.step_destroy(None)


def test_capture_errors(M):
    L = M._lib
    with pytest.raises(ValueError, match="default stream"):
        L.step_capture_begin(None)
    with pytest.raises(ValueError, match="null graph"):
        L.step_launch(None, torch.cuda.current_stream().cuda_stream)
    s = torch.cuda.Stream()
    with pytest.raises(RuntimeError):                   # ending a capture that was never begun is a CUDA error
        L.step_capture_end(s.cuda_stream, ctypes.byref(ctypes.c_void_p()), None)
    torch.cuda.synchronize()

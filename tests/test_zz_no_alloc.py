"""lmsfc_colsum_static: the allocation-free variant of lmsfc_colsum (include/b200lmsfc.h), called through the C ABI and
checked against the existing entry point on the same inputs, on consecutive calls that share the static scratch.

Collected last on purpose.  The entry point has no counterpart in the reference.
Tolerance: fp32 column sums of bf16 data, rel 2e-4 of the largest column (the fp32 tolerance of test_gpu_kernels.py);
both entry points run the same kernel on the same grid, so the sums are in fact expected to agree bit for bit.
"""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
L = None


def _lib():
    global L
    if L is None:
        L = importlib.import_module("l-msfc-v2-training_b200._lib")
    return L


def _both(g, C):
    """(existing entry point, static entry point) column sums of g [P][C] bf16, each into a zeroed fp32 [C]."""
    lib = _lib()
    P = g.numel() // C
    want = torch.zeros(C, device=DEV, dtype=torch.float32)
    got = torch.zeros(C, device=DEV, dtype=torch.float32)
    lib.colsum(lib.ptr(g), P, C, lib.ptr(want), lib.stream())
    lib.colsum_static(lib.ptr(g), P, C, lib.ptr(got), lib.stream())
    return want, got


def _check(want, got, g, C):
    ref = g.float().reshape(-1, C).sum(0)
    tol = 2e-4 * float(ref.abs().max().clamp_min(1.0))
    assert float((got - want).abs().max()) <= tol, (float((got - want).abs().max()), tol)
    assert float((got - ref).abs().max()) <= 2e-3 * float(g.float().abs().reshape(-1, C).sum(0).max()), "static colsum is off"


# consecutive calls of different shapes: the second reads scratch rows the first one left behind
CASES = [(4096, 256), (777, 1024), (16 * 64 * 64, 256), (3, 64), (2500, 1024), (4096, 256)]


def test_static_colsum_matches_colsum_on_consecutive_calls():
    torch.manual_seed(7)
    for i, (P, C) in enumerate(CASES):
        g = (torch.randn(P, C, device=DEV) * (1.0 + i)).to(torch.bfloat16)
        want, got = _both(g, C)
        torch.cuda.synchronize()
        _check(want, got, g, C)
    # twice in a row on the SAME stream without a sync in between, accumulating into one output (out += ...)
    P, C = 5000, 512
    g1 = torch.randn(P, C, device=DEV).to(torch.bfloat16)
    g2 = torch.randn(P, C, device=DEV).to(torch.bfloat16)
    lib = _lib()
    want = torch.zeros(C, device=DEV)
    got = torch.zeros(C, device=DEV)
    for g in (g1, g2):
        lib.colsum(lib.ptr(g), P, C, lib.ptr(want), lib.stream())
    for g in (g1, g2):
        lib.colsum_static(lib.ptr(g), P, C, lib.ptr(got), lib.stream())
    torch.cuda.synchronize()
    _check(want, got, torch.cat([g1, g2]), C)


def test_static_colsum_allocates_nothing():
    lib = _lib()
    P, C = 4096, 1024
    g = torch.randn(P, C, device=DEV).to(torch.bfloat16)
    out = torch.zeros(C, device=DEV)
    lib.colsum_static(lib.ptr(g), P, C, lib.ptr(out), lib.stream())      # set-up call (symbol addresses, module load)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    a0 = torch.cuda.memory_allocated()
    for _ in range(3):
        lib.colsum_static(lib.ptr(g), P, C, lib.ptr(out), lib.stream())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_allocated() == a0
    # under stream capture (cudaMalloc would fail the capture in the default "global" error mode)
    out2 = torch.zeros(C, device=DEV)
    want = torch.zeros(C, device=DEV)
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(graph, stream=s):
            lib.colsum_static(lib.ptr(g), P, C, lib.ptr(out2), lib.stream())
    torch.cuda.current_stream().wait_stream(s)
    graph.replay()
    lib.colsum(lib.ptr(g), P, C, lib.ptr(want), lib.stream())
    torch.cuda.synchronize()
    _check(want, out2, g, C)


def test_static_colsum_rejects_sizes_beyond_the_static_scratch():
    lib = _lib()
    C = 1024 + 8
    g = torch.zeros(8, C, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros(C, device=DEV)
    with pytest.raises(ValueError, match="exceeds the static scratch"):
        lib.colsum_static(lib.ptr(g), 8, C, lib.ptr(out), lib.stream())
    with pytest.raises(ValueError, match="multiple of 8"):
        lib.colsum_static(lib.ptr(g), 8, 100, lib.ptr(out), lib.stream())
    assert float(out.abs().max()) == 0.0

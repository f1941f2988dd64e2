"""The data-parallel exchange kernels (csrc/exchange.cu) on ONE GPU: N simulated ranks = N buffer sets on the same
device, each rank's kernels on its own stream, so the flag protocol, the fixed-order reduction, the norm exchange and
the sharded Adam + push are exercised exactly as between GPUs (tests/test_gpu_dp.py runs the real thing on 2 GPUs).
Reference: torch -- mean of the ranks' gradients, clip_grad_norm_(1.0), one Adam step (examples/train.py:100-103).
"""
import ctypes
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
PKG = "l-msfc-v2-training_b200"


def _peers(L, world, rank, bufs, tick):
    p = L.XchgPeers()
    p.world, p.rank = world, rank
    for k in range(world):
        p.grads[k] = bufs[k]["grads"].data_ptr()
        p.params[k] = bufs[k]["params"].data_ptr()
        p.shadow[k] = bufs[k]["shadow"].data_ptr()
        p.flags[k] = bufs[k]["flags"].data_ptr()
        p.norm_all[k] = bufs[k]["norm_all"].data_ptr()
    p.tick = tick.data_ptr()
    return p


@pytest.mark.parametrize("world", [2, 4, 8])
def test_exchange_on_simulated_ranks(world):
    L = importlib.import_module(f"{PKG}._lib")
    parallel = importlib.import_module(f"{PKG}.parallel")
    n = 64 * 1000                      # arena size (floats)
    fp32_end = 64 * 100
    units = [(0, 64 * 300), (64 * 300, n)]
    g = torch.Generator().manual_seed(world)
    p0 = torch.randn(n, generator=g).to(DEV)
    bufs, ticks, opt = [], [], []
    for r in range(world):
        bufs.append(dict(grads=(torch.randn(n, generator=g) * 0.01).to(DEV), params=p0.clone(),
                         shadow=torch.zeros(n, dtype=torch.bfloat16, device=DEV),
                         flags=torch.zeros(L.XCHG_SLOTS * L.MAX_RANKS, dtype=torch.int64, device=DEV),
                         norm_all=torch.zeros(L.MAX_RANKS, device=DEV)))
        ticks.append(torch.zeros(1, dtype=torch.int64, device=DEV))
        opt.append(dict(m=torch.zeros(n, device=DEV), v=torch.zeros(n, device=DEV), step=torch.zeros(1, dtype=torch.int32, device=DEV),
                        scratch=torch.zeros(8, device=DEV), sumsq=torch.zeros(1, device=DEV), parts=torch.zeros(len(units), device=DEV)))
    g_mean = torch.stack([b["grads"] for b in bufs]).mean(0)
    streams = [torch.cuda.Stream() for _ in range(world)]
    # one process plays every rank: nothing may block the host while a rank's kernel spins, so load all kernels first
    L.xchg_init()
    dummy = dict(step=torch.zeros(1, dtype=torch.int32, device=DEV), scratch=torch.zeros(8, device=DEV))
    L.adam_prep(1e-2, None, 0.9, 0.999, 0.0, None, dummy["step"].data_ptr(), dummy["scratch"].data_ptr(), None, 1, None, L.stream())
    torch.cuda.synchronize()
    for step in range(2):
        if step == 1:                  # second step: fresh gradients, the protocol's epochs move on
            for r in range(world):
                bufs[r]["grads"].copy_((torch.randn(n, generator=g) * 0.01).to(DEV))
            g_mean2 = torch.stack([b["grads"] for b in bufs]).mean(0)
            torch.cuda.synchronize()
        for r in range(world):
            peers = _peers(L, world, r, bufs, ticks[r])
            o = opt[r]
            with torch.cuda.stream(streams[r]):
                st = L.stream()
                for u, (lo, hi) in enumerate(units):
                    a, b = parallel.owned_range(lo, hi, r, world)
                    L.xchg_reduce(peers, u, a, b, o["parts"][u:].data_ptr(), 4, st)
                L.xchg_norm_publish(peers, o["parts"].data_ptr(), len(units), 8, None, st)
                L.xchg_norm_wait(peers, 8, o["sumsq"].data_ptr(), st)
                L.adam_prep(1e-2, None, 0.9, 0.999, 1.0, o["sumsq"].data_ptr(), o["step"].data_ptr(), o["scratch"].data_ptr(),
                            None, 1, None, st)
                for u, (lo, hi) in enumerate(units):
                    a, b = parallel.owned_range(lo, hi, r, world)
                    L.xchg_adam_push(peers, 16 + u, lo, hi, a, b, fp32_end, o["m"].data_ptr(), o["v"].data_ptr(), 0.9, 0.999,
                                     1e-8, o["scratch"].data_ptr(), 4, st)
                for u in range(len(units)):
                    L.xchg_wait(peers, 16 + u, st)
        torch.cuda.synchronize()
    # ---- reference: two clipped Adam steps on the mean gradients
    pr = torch.nn.Parameter(p0.clone())
    ropt = torch.optim.Adam([pr], lr=1e-2)
    for gm in (g_mean, g_mean2):
        pr.grad = gm.clone()
        torch.nn.utils.clip_grad_norm_([pr], 1.0)
        ropt.step()
    want = pr.detach()
    full = torch.empty(n, device=DEV)              # the owners' fp32 master values, gathered
    for r in range(world):
        for lo, hi in units:
            a, b = parallel.owned_range(lo, hi, r, world)
            full[a:b] = bufs[r]["params"][a:b]
    torch.testing.assert_close(full, want, rtol=1e-5, atol=1e-6)
    for r in range(world):
        assert torch.equal(bufs[r]["shadow"], bufs[0]["shadow"])                       # replicas bit-identical
        assert torch.equal(bufs[r]["params"][:fp32_end], bufs[0]["params"][:fp32_end])
        torch.testing.assert_close(bufs[r]["params"][:fp32_end], want[:fp32_end], rtol=1e-5, atol=1e-6)
        assert torch.equal(bufs[r]["shadow"], full.to(torch.bfloat16))
        assert float(bufs[r]["grads"].abs().max()) == 0.0                              # cleared for the next backward
        assert float(opt[r]["scratch"][0]) == float(opt[0]["scratch"][0])              # one clip coefficient
        assert int(ticks[r]) == 2
    norm = float(opt[0]["scratch"][3])
    assert abs(norm - float(g_mean2.norm())) < 1e-4 * norm

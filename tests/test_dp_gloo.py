"""Host logic of the data-parallel gradient exchange (parallel.BucketedGradReducer) on CPU:
world_size 2 over gloo.  The CUDA path differs only in where the all-reduce is enqueued (a side
stream behind an event, NCCL AVG) -- bucket layout, readiness tracking, launch order, the
"quantiles are never reduced" rule and the averaged result are what is checked here.
"""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

PKG = "l-msfc-v2-training_b200"


class _FakeArena:
    """The slice of model.ParamArena the reducer uses (flat grads + offsets), on host memory."""
    ALIGN = 64

    def __init__(self, model):
        named = list(model.named_parameters())
        self.net = [(n, p) for n, p in named if not n.endswith(".quantiles")]
        self.aux = [(n, p) for n, p in named if n.endswith(".quantiles")]
        self.offsets, off = {}, 0
        for n, p in self.net:
            self.offsets[n] = off
            off += (p.numel() + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.net_grads = torch.zeros(off)
        self.aux_grads = torch.zeros(sum(p.numel() for _, p in self.aux))
        self.device = torch.device("cpu")

    def names(self, group="net"):
        return [n for n, _ in (self.net if group == "net" else self.aux)]


class _Layer(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.weight = nn.Parameter(torch.zeros(n))
        self.bias = nn.Parameter(torch.zeros(8))
        self._state = type("S", (), {"grad_ready": None})()


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.enc = _Layer(3000)
        self.entropy_bottleneck = nn.Module()
        self.entropy_bottleneck._matrix0 = nn.Parameter(torch.zeros(100))
        self.entropy_bottleneck.quantiles = nn.Parameter(torch.zeros(12))
        self.dec1 = _Layer(5000)
        self.dec2 = _Layer(2000)
        self.arena = None


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        par = importlib.import_module(PKG + ".parallel")
        net = _Net()
        net.arena = _FakeArena(net)
        # replicas built from different seeds: construction must leave every rank with rank 0's parameters
        # (what DistributedDataParallel does; the reference's nn.DataParallel replicates from GPU 0 every step)
        net.arena.net_params = torch.full((64,), float(rank + 1))
        net.arena.aux_params = torch.full((12,), float(10 * rank + 5))
        red = par.BucketedGradReducer(net, bucket_mb=2048 * 4 / (1 << 20))      # 2048-float buckets
        ar = net.arena
        assert torch.all(ar.net_params == 1.0) and torch.all(ar.aux_params == 5.0)
        # buckets tile the arena, back to front, without gaps
        spans = sorted((b["lo"], b["hi"]) for b in red.buckets)
        assert spans[0][0] == 0 and spans[-1][1] == ar.net_grads.numel()
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert red.buckets[0]["hi"] == ar.net_grads.numel()                     # first bucket = end of arena (decoder)
        assert "entropy_bottleneck.quantiles" not in red.bucket_of              # aux parameters are never reduced
        assert len(red.buckets) >= 2
        # every hooked layer reports through the same callback
        assert all(m._state.grad_ready is not None for m in (net.enc, net.dec1, net.dec2))
        for step in range(2):
            ar.net_grads.fill_(float(rank + 1 + step))
            ar.aux_grads.fill_(float(rank + 7))
            red.begin()
            # "backward": decoder first; a bucket launches as soon as its last gradient is ready
            net.dec2._state.grad_ready([net.dec2.weight, net.dec2.bias])
            net.dec1._state.grad_ready([net.dec1.weight])
            launched_mid = list(red.launched)
            net.dec1._state.grad_ready([net.dec1.bias])
            red.mark_ready(red.eb_names)
            net.enc._state.grad_ready([net.enc.weight, net.enc.bias])
            red.finish()
            assert sorted(red.launched) == list(range(len(red.buckets)))
            assert launched_mid == [0] and red.launched[:2] == [0, 1]   # overlap: a bucket goes as soon as it is full
            want = sum(r + 1 + step for r in range(world)) / world
            assert torch.allclose(ar.net_grads, torch.full_like(ar.net_grads, want))
            assert torch.all(ar.aux_grads == rank + 7)                          # untouched
        # a bucket whose parameters never reported is flushed by finish()
        ar.net_grads.fill_(float(rank))
        red.begin()
        red.finish()
        assert torch.allclose(ar.net_grads, torch.full_like(ar.net_grads, sum(range(world)) / world))
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_bucketed_reducer_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=100) for _ in procs]
    for p in procs:
        p.join(20)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_reducer_single_process_is_a_noop():
    par = importlib.import_module(PKG + ".parallel")
    net = _Net()
    net.arena = _FakeArena(net)
    red = par.BucketedGradReducer(net, bucket_mb=0.001)
    net.arena.net_grads.fill_(3.0)
    red.begin()
    red.finish()
    assert torch.all(net.arena.net_grads == 3.0)
    assert sorted(red.launched) == list(range(len(red.buckets)))


def test_owned_ranges_partition_every_unit():
    """parallel.owned_range: the N owners' parts tile a unit exactly, on 64-float boundaries, identically on every rank
    (the reduce-scatter / sharded Adam / all-gather kernels of csrc/exchange.cu rely on it)."""
    par = importlib.import_module(PKG + ".parallel")
    for lo, hi in [(0, 2052352), (2052352, 5329152), (19616000, 27939072), (64, 64 * 3), (0, 64), (128, 128)]:
        for world in (1, 2, 3, 4, 8):
            parts = [par.owned_range(lo, hi, r, world) for r in range(world)]
            assert parts[0][0] == lo and parts[-1][1] == hi
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            assert all(a % 64 == 0 and b % 4 == 0 and b >= a for a, b in parts)
            if hi - lo >= 64 * world * 8:
                sizes = [b - a for a, b in parts]
                assert max(sizes) - min(sizes) <= 64 + (hi - lo) % 64

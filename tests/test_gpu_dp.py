"""Data-parallel train step on 2 GPUs over NCCL (skipped on a single-GPU box; run with `gpurun --gpus 2`).

* eager DP step with injected noise == the single-process step on the concatenated batch (gradient AVG over ranks is
  the gradient of the global-batch mean loss; examples/train.py:64-71 wraps the model in DataParallel for the same effect);
* the captured, pipelined DP step (NCCL kernels inside the CUDA graph, 3 weight-gradient streams) leaves bit-identical
  weights on both ranks and trains.
"""
import importlib
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu
PKG = "l-msfc-v2-training_b200"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    from oracle import lmsfc as OL                      # synthetic feature generator only
    spec_m, model_m, train_m = (importlib.import_module(f"{PKG}.{m}") for m in ("spec", "model", "train"))
    spec = spec_m.DEFAULT_SPEC
    B = 2
    feats_all = OL.synth_features(spec, world * B, 32, 32, seed=7)
    feats_all = {k: v.to(torch.bfloat16).float() for k, v in feats_all.items()}
    g = torch.Generator().manual_seed(3)
    nz_all = torch.rand((world * B, 256, 1, 1), generator=g) - 0.5
    ny_all = torch.rand((world * B, 256, 2, 2), generator=g) - 0.5
    sl = slice(rank * B, (rank + 1) * B)
    feats = {k: v[sl].to(dev) for k, v in feats_all.items()}

    # ---- eager DP, injected noise.  The ranks build their replicas from DIFFERENT seeds: construction must leave every
    # rank with rank 0's weights (the reducer used to rely on identical seeding)
    torch.manual_seed(21 + 5 * rank)
    net = model_m.LMSFC(spec).to(dev)
    ts = train_m.TrainStep(net, lmbda=0.013, distributed=True)
    start = net.arena.net_params.clone()
    dist.broadcast(start, src=0)
    out[f"start_equal_{rank}"] = bool(torch.equal(start, net.arena.net_params))
    out[f"noise_seed_{rank}"] = int(net.entropy_bottleneck._rng(dev).state[0])
    for _ in range(2):
        ts.step(feats, nz_all[sl].to(dev), ny_all[sl].to(dev))
    ts.consolidate()                                    # sharded exchange: gather the owners' fp32 masters
    torch.cuda.synchronize()
    dp_params = net.arena.net_params.clone()
    # ---- the same two steps in one process on the concatenated batch (rank 0 only)
    if rank == 0:
        torch.manual_seed(21)
        ref = model_m.LMSFC(spec).to(dev)
        ts_ref = train_m.TrainStep(ref, lmbda=0.013, distributed=False)
        fa = {k: v.to(dev) for k, v in feats_all.items()}
        for _ in range(2):
            ts_ref.step(fa, nz_all.to(dev), ny_all.to(dev))
        torch.cuda.synchronize()
        out["eager_rel"] = float((dp_params - ref.arena.net_params).norm() / ref.arena.net_params.norm())
        out["eager_step_rel"] = float((dp_params - ref.arena.net_params).norm() /
                                      (ref.arena.net_params - _init_params(model_m, spec, dev)).norm())
    # ---- captured, pipelined DP step
    ts.capture(feats, warmup=1, pipelined=True)
    losses = [float(ts.replay()["loss"]) for _ in range(6)]
    ts.consolidate()
    torch.cuda.synchronize()
    out[f"shadow_equal_{rank}"] = bool(torch.equal(net.arena.net_bf16, _bcast0(dist, net.arena.net_bf16)))
    mine = net.arena.net_params.clone()
    other = mine.clone()
    dist.broadcast(other, src=0)
    out[f"graph_equal_{rank}"] = bool(torch.equal(mine, other))
    out[f"losses_{rank}"] = losses
    dist.barrier()
    torch.cuda.synchronize()
    ts.release()
    del ts
    torch.cuda.synchronize()
    dist.destroy_process_group()


def _bcast0(dist, t):
    o = t.clone()
    dist.broadcast(o, src=0)
    return o


def _init_params(model_m, spec, dev):
    torch.manual_seed(21)
    return model_m.LMSFC(spec).to(dev).build_arena().net_params.clone()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_dp_two_gpus_matches_single_process_and_ranks_agree():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
    for p in procs:
        if p.is_alive():
            p.kill()
        assert p.exitcode == 0, p.exitcode
    res = dict(out)
    print(res)
    # distance between the DP and the single-process weights, relative to the distance the two optimiser steps
    # travelled (measured 0.023 on 2 B200s: bf16 GEMM summation order flips Adam's sign-like first updates only
    # where the gradient is ~0)
    assert res["eager_step_rel"] < 0.1, res
    assert res["graph_equal_0"] and res["graph_equal_1"], res
    assert res["shadow_equal_0"] and res["shadow_equal_1"], res
    assert res["start_equal_0"] and res["start_equal_1"], res
    assert res["noise_seed_0"] != res["noise_seed_1"], res      # independent quantisation noise per replica
    l0, l1 = res["losses_0"], res["losses_1"]
    assert all(l == l and l < 1e6 for l in l0 + l1)
    assert l0[-1] < l0[0] * 1.5 and len(set(l0)) > 3

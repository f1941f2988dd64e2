"""BASELINE configs[2]: the lambda sweep of 6 rate points (compressai/docs/source/zoo.rst:68), each a data-parallel
training run of K captured steps on synthetic batch-16/GPU feature crops; run sequentially (what BASELINE.json asks).
  python profiles/lambda_sweep.py [K]                                   # one GPU
  python -m torch.distributed.run --nproc-per-node 8 ... profiles/lambda_sweep.py [K]
Prints one JSON line per lambda (rank 0): ms/step, feature-sets/s, loss / bpp / mse at the first and the last step."""
import importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
spec_m, model_m, train_m, data_m = (importlib.import_module(f"{PKG}.{m}") for m in ("spec", "model", "train", "data"))
K = int(sys.argv[1]) if len(sys.argv) > 1 else 300
sp = spec_m.DEFAULT_SPEC
feats = [{k: v.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).to(dev)
          for k, v in data_m.synth_features(sp, 16, 64, 64, seed=1234 + rank * 100 + i).items()} for i in range(4)]
for lmbda in spec_m.LAMBDAS:
    torch.manual_seed(1234)
    net = model_m.LMSFC(sp).to(dev)
    ts = train_m.TrainStep(net, lmbda=lmbda, distributed=world > 1)
    first = {k: float(v) for k, v in ts.step(feats[0]).items()}
    ts.capture(feats[0], warmup=1, pipelined=True)
    for i in range(10):
        ts.replay(feats[i % 4])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(K):
        out = ts.replay(feats[i % 4])
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / K], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    last = out["scalars"].tolist()
    ts.flush()
    if rank == 0:
        print(json.dumps({"lambda": lmbda, "n_gpus": world, "steps": K, "ms_per_step": float(ms), "feature_sets_per_s": world * 16 / (float(ms) * 1e-3),
                          "first": first, "last": dict(zip(("loss", "bpp_loss", "mse_loss", "aux_loss"), last))}), flush=True)
    if world > 1:
        dist.barrier(); torch.cuda.synchronize()
    ts.release(); del ts, net
    torch.cuda.empty_cache()
if world > 1:
    dist.destroy_process_group()

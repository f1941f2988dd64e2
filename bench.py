"""bench.py -- L-MSFC train-step throughput (feature-sets/s) on N B200s; see DESIGN.md §Measurement.

  python bench.py --gpus N --steps K --warmup W            # this framework (CUDA kernels via C ABI)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

Workload = BASELINE.json configs[1]: batch 16 synthetic P2-P5 feature crops (256x256 input ->
P2 64x64 ... P5 8x8, 256 channels), bf16, one lambda, per GPU (weak scaling).
A step = zero_grad -> forward -> RD loss -> backward -> [all-reduce] -> clip + Adam -> aux Adam.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"
BATCH, H2, W2, LMBDA = 16, 64, 64, 0.0130
METRIC, UNIT = "L-MSFC train feature-sets/sec", "feature-sets/s"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop = index, [], False

    def run(self):
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in o.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        mx = max((int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's own CPU implementation of the step: its modules as restated in oracle/
    (PyTorch fp32 eager, all host threads), bounded sample of the same workload."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import lmsfc as OL
    sp = _load_spec()
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    bs = 2                                   # bounded sample: 2 of the 16 feature-sets per step
    net = OL.OracleLMSFC(sp)
    opt, aux = OL.configure_optimizers(net)
    feats = OL.synth_features(sp, bs, H2, W2)
    for _ in range(max(1, min(args.warmup, 1))):
        OL.train_step(net, opt, aux, feats, LMBDA)
    steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(steps):
        OL.train_step(net, opt, aux, feats, LMBDA)
    dt = time.perf_counter() - t0
    v = bs * steps / dt
    sample = f"{steps} steps of batch {bs} (of {BATCH}) 256x256-crop feature-sets, fp32, torch {torch.__version__} CPU"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": 1, "ms_per_step": 1e3 * dt / steps * BATCH / bs, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "L-MSFC training, batch 16 synthetic P2-P5 crops (256x256 input), one lambda",
                   "lambda": LMBDA, "note": "CPU path, rank 0 only"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def _load_spec():
    return importlib.import_module(PKG + ".spec").DEFAULT_SPEC


def cpu_baseline(sp):
    import torch
    from oracle import lmsfc as OL
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    bs = 2
    net = OL.OracleLMSFC(sp)
    opt, aux = OL.configure_optimizers(net)
    feats = OL.synth_features(sp, bs, H2, W2)
    OL.train_step(net, opt, aux, feats, LMBDA)
    t0 = time.perf_counter()
    n = 2
    for _ in range(n):
        OL.train_step(net, opt, aux, feats, LMBDA)
    dt = time.perf_counter() - t0
    return {"value": bs * n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} steps of batch {bs} (of {BATCH}) feature-sets, oracle port (torch {torch.__version__} fp32 CPU)"}


# ------------------------------------------------------------------------------- kernel rooflines
def kernel_rooflines(torch, L, layers, dev):
    """Dominant kernels timed live with CUDA events on the launching stream (each launch bracketed,
    averaged): conv2-shaped forward implicit GEMM and its weight-gradient GEMM.
    Algorithmic FLOPs per launch = 2 * M * N * K (SURVEY.md §8d / DESIGN.md)."""
    pk, which = peaks()
    conv = layers.Conv(256, 256, 5, 2).to(dev)
    x = torch.randn(BATCH, 256, H2, W2, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x.requires_grad_()
    res = {}
    stream = torch.cuda.current_stream()
    y = conv(x)
    g = torch.randn_like(y)
    y.backward(g)                                   # warm-up of fwd, dgrad, wgrad
    M, N, K = BATCH * (H2 // 2) * (W2 // 2), 256, 25 * 256
    flops = 2.0 * M * N * K

    def timeit(fn, n=10, reps=20):
        """Average launch duration: n launches recorded into a CUDA graph (the host side of one eager call -- ctypes,
        tensor-map encode -- is longer than the kernel), replayed `reps` times between CUDA events on the launching
        stream; the median replay / n."""
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(stream)
        with torch.cuda.stream(side):
            for _ in range(3):
                fn()
        stream.wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n):
                fn()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); g.replay(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) / n)
        ts.sort()
        return ts[len(ts) // 2]

    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
    except Exception:
        traffic = {}

    st = conv._state
    d = st.desc(BATCH, H2, W2, H2 // 2, W2 // 2)
    yb = torch.empty_like(y)
    P = L.ptr
    ms_f = timeit(lambda: L.conv_forward(d, P(x), None, P(st.w_direct), P(conv.bias), 0, P(yb), None, None, L.stream()))
    gw = torch.zeros((256, 25, 256), device=dev)
    gb = g.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ms_w = timeit(lambda: L.conv_wgrad(d, P(x), None, P(gb), P(gw), None, L.stream()))
    for name, ms in (("igemm_fwd_kernel<256>", ms_f), ("wgrad_kernel", ms_w)):
        a = flops / (ms * 1e-3) / 1e12
        res[name] = {"bound": "tensor", "achieved": a, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                     "frac": a / pk["bf16_tflops"], "traffic": (traffic.get(name) or {}).get("bytes"), "ms_per_launch": ms,
                     "flops_per_launch": flops, "peak_source": f"bf16_tflops (burst) of {which}",
                     "shape": f"M={M} N={N} K={K} (g_a.conv2 at batch 16)"}
    return res


# ------------------------------------------------------------------------------- this framework
def run_b200(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this framework has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = importlib.import_module(PKG + "._lib")
    model_m = importlib.import_module(PKG + ".model")
    train_m = importlib.import_module(PKG + ".train")
    layers = importlib.import_module(PKG + ".layers")
    data_m = importlib.import_module(PKG + ".data")
    sp = model_m.DEFAULT_SPEC
    torch.manual_seed(1234)
    net = model_m.LMSFC(sp).to(dev)
    ts = train_m.TrainStep(net, lmbda=LMBDA, distributed=world > 1, bucket_mb=float(os.environ.get("LMSFC_BUCKET_MB", "64")))
    nbuf = 4
    host = []
    for i in range(nbuf):                    # pinned bf16 NHWC feature shards, one per buffered step
        f = data_m.synth_features(sp, BATCH, H2, W2, seed=1234 + rank * 100 + i)
        host.append({k: v.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).pin_memory() for k, v in f.items()})
    feats = {k: v.to(dev, non_blocking=True) for k, v in host[0].items()}
    h2d = sum(v.numel() * 2 for v in host[0].values())
    def note(msg):
        if args.verbose:
            print(f"[bench rank {rank}] {msg}", file=sys.stderr, flush=True)

    note("model built")
    ts.step(feats)                           # eager once (lazy init), then capture the graph
    torch.cuda.synchronize()
    note("eager step done")
    ts.capture(feats, warmup=1, pipelined=not args.sequential, n_buffers=2)
    launches_per_step = ts.launches_per_step               # launches of this library recorded into the graph
    torch.cuda.synchronize()
    note("graph captured")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value")
    for _ in range(max(args.warmup, 3)):
        ts.replay()
    torch.cuda.synchronize()
    note("warm replays done")
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        ts.replay()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # ---- end to end: pinned host features -> H2D (copy stream, double buffered) -> step -> D2H loss
    copy_stream = torch.cuda.Stream(device=dev)
    dbuf = [ts.input_buffers(0), ts.input_buffers(1)]     # H2D lands directly in the buffers the two graphs read
    loss_host = torch.zeros(args.steps, 4).pin_memory()
    main = torch.cuda.current_stream()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    for i in range(args.steps + 1):
        if i < args.steps:                   # stage step i's inputs
            b = i % 2
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[b])
                for k, v in host[i % nbuf].items():
                    dbuf[b][k].copy_(v, non_blocking=True)
                ready[b].record(copy_stream)
        if i >= 1:                           # run step i-1
            b = (i - 1) % 2
            main.wait_event(ready[b])
            out = ts.replay(buffer=b)
            consumed[b].record(main)
            loss_host[i - 1].copy_(out["scalars"], non_blocking=True)
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    ts.flush()                               # pipelined order: apply the update of the last backward pass
    if sampler:
        sampler.stop = True
        sampler.join()
    t = torch.tensor([ms, ms_e2e], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    if rank == 0:
        roof = kernel_rooflines(torch, L, layers, dev)
        dom = max(roof, key=lambda k: roof[k]["ms_per_launch"])
        pk, which = peaks()
        flops_step = sp.flops_train(H2, W2) * BATCH
        line = {
            "metric": METRIC, "value": world * BATCH * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "L-MSFC training, batch 16/GPU synthetic P2-P5 feature crops (256x256 input), bf16, one lambda",
                       "lambda": LMBDA, "spec": "N=M=F=256 (spec.py)", "cuda_graph": True,
                       "step_order": "sequential" if args.sequential else
                                     "pipelined: each step = [Adam of the previous backward || forward] -> backward (train.py)",
                       "l2": "per-step working set (weights+Adam state+activations ~0.9 GB) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}"},
            "e2e": {"value": world * BATCH * args.steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 16, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step,
            "step_tensor_tflops": flops_step / (ms / args.steps * 1e-3) / 1e12,
            "step_tensor_frac_of_sustained": flops_step / (ms / args.steps * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
            "roofline": dict(roof[dom], kernel=dom), "kernels": roof,
            "clocks": sampler.summary() if sampler else None,
            "final_loss": float(loss_host[-1, 0]),
        }
        if world == 1:                       # the CPU baseline is measured at N=1 only (rank 0, idle host cores)
            try:
                line["cpu_baseline"] = cpu_baseline(sp)
            except Exception as e:  # the baseline is reported, never a dependency
                line["cpu_baseline"] = {"value": None, "error": str(e)}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    if world > 1:
        # Tear down in order: the captured graph holds NCCL kernels, and destroying the communicator under a live
        # graph hung (measured, 2 GPUs).  Drop the graph first; exit hard if the teardown still does not return.
        dist.barrier()
        torch.cuda.synchronize()
        ts.release()
        del ts
        torch.cuda.synchronize()
        killer = threading.Timer(20.0, lambda: os._exit(0))
        killer.daemon = True
        killer.start()
        dist.destroy_process_group()
        killer.cancel()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--verbose", action="store_true", help="progress notes on stderr")
    ap.add_argument("--sequential", action="store_true",
                    help="capture the eager order (update at the end of the step) instead of the rotated, pipelined one")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

"""Stream-level timeline of ONE captured training step (CUPTI activity records through torch.profiler).
  python profiles/trace_step.py [out.txt]
Prints every kernel of the last graph replay (start us, duration us, stream, name), the busy time per
stream, and the time no kernel at all was running.  A diagnosis tool: numbers taken under the profiler
are never bench values."""
import importlib, json, os, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"
model_m = importlib.import_module(PKG + ".model")
train_m = importlib.import_module(PKG + ".train")

world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
if world > 1:                          # torchrun: every rank runs the step, rank 0 writes its own timeline
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
sp = model_m.DEFAULT_SPEC
torch.manual_seed(1234)
net = model_m.LMSFC(sp).to(dev)
ts = train_m.TrainStep(net, lmbda=0.013, distributed=world > 1)
f = importlib.import_module(PKG + '.data').synth_features(sp, 16, 64, 64, seed=1234)
feats = {k: v.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).to(dev) for k, v in f.items()}
ts.step(feats)
ts.capture(feats, warmup=1, pipelined=os.environ.get("LMSFC_SEQUENTIAL", "0") != "1")
for _ in range(5):
    ts.replay()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        ts.replay()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), f"lmsfc_trace_{rank}.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])
n = len(ev) // 3
ev = ev[2 * n:]                        # the last replay
t0 = ev[0]["ts"]
# rank 0 writes out.txt, rank k > 0 out.txt.rank<k> (absolute times are per-rank clocks: compare durations and order)
out = open(sys.argv[1] + (f".rank{rank}" if rank else ""), "w") if len(sys.argv) > 1 else (sys.stdout if rank == 0 else open(os.devnull, "w"))
streams = {}
for e in ev:
    s = e["args"].get("stream", -1)
    streams.setdefault(s, len(streams))
print(f"# {len(ev)} device activities in one step; streams {streams}", file=out)
busy = {s: 0.0 for s in streams}
for e in ev:
    s = e["args"].get("stream", -1)
    busy[s] += e["dur"]
    g = e["args"].get("grid")
    gs = "x".join(str(v) for v in g) if g else "-"
    nm = e["name"].replace("lmsfc::", "").replace("void ", "")
    print(f"{e['ts'] - t0:9.2f} {e['dur']:8.2f}  s{streams[s]}  {gs:>10s}  {nm[:80]}", file=out)
end = max(e["ts"] + e["dur"] for e in ev)
# union of busy intervals
iv = sorted((e["ts"], e["ts"] + e["dur"]) for e in ev)
cov, cur_s, cur_e = 0.0, iv[0][0], iv[0][1]
for a, b in iv[1:]:
    if a > cur_e:
        cov += cur_e - cur_s
        cur_s, cur_e = a, b
    else:
        cur_e = max(cur_e, b)
cov += cur_e - cur_s
print(f"# span {end - t0:.1f} us; any-kernel-running {cov:.1f} us; idle {end - t0 - cov:.1f} us", file=out)
for s, i in streams.items():
    print(f"# stream s{i}: busy {busy[s]:.1f} us", file=out)

out.flush()
if world > 1:
    dist.barrier(); torch.cuda.synchronize(); ts.release(); del ts; dist.destroy_process_group()

"""Why is the end-to-end loop slower than graph replays?  torchrun ... profiles/tools/e2e_probe.py  (rank 0 prints)."""
import importlib, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1: dist.init_process_group("nccl", device_id=dev)
model_m, train_m, data_m = (importlib.import_module(f"{PKG}.{m}") for m in ("model", "train", "data"))
sp = model_m.DEFAULT_SPEC
torch.manual_seed(1234)
net = model_m.LMSFC(sp).to(dev)
if os.environ.get("LMSFC_PROBE_SYMM") == "1":      # NCCL exchange, but the arena in peer-mapped memory
    par = importlib.import_module(f"{PKG}.parallel")
    net.build_arena(alloc=par.symmetric_alloc())
ts = train_m.TrainStep(net, lmbda=0.013, distributed=world > 1)
host = [{k: v.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).pin_memory()
         for k, v in data_m.synth_features(sp, 16, 64, 64, seed=1234 + rank * 100 + i).items()} for i in range(4)]
if os.environ.get("LMSFC_PROBE_D2D") == "1":       # "host" buffers on the device: no PCIe traffic
    host = [{k: v.to(dev) for k, v in h.items()} for h in host]
feats = {k: v.to(dev) for k, v in host[0].items()}
ts.step(feats); ts.capture(feats, warmup=1, pipelined=True, n_buffers=2)
for _ in range(10): ts.replay()
def timed(fn, n=200):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); fn(n); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
res = {}
res["graph0 only"] = timed(lambda n: [ts.replay(buffer=0) for _ in range(n)])
res["alternating graphs"] = timed(lambda n: [ts.replay(buffer=i % 2) for i in range(n)])
cs = torch.cuda.Stream(device=dev)
dbuf = [ts.input_buffers(0), ts.input_buffers(1)]
def h2d_only(n):
    with torch.cuda.stream(cs):
        for i in range(n):
            for k, v in host[i % 4].items(): dbuf[i % 2][k].copy_(v, non_blocking=True)
    torch.cuda.current_stream().wait_stream(cs)
res["h2d only"] = timed(h2d_only)
def unsynced(n):          # copies and replays with no dependency between them
    with torch.cuda.stream(cs):
        for i in range(n):
            for k, v in host[i % 4].items(): dbuf[i % 2][k].copy_(v, non_blocking=True)
    for i in range(n): ts.replay(buffer=0)
    torch.cuda.current_stream().wait_stream(cs)
res["h2d beside replays (no deps)"] = timed(unsynced)
def e2e(n, d2h=True):
    main = torch.cuda.current_stream()
    ready = [torch.cuda.Event(), torch.cuda.Event()]; consumed = [torch.cuda.Event(), torch.cuda.Event()]
    lh = torch.zeros(n, 4).pin_memory()
    for i in range(n + 1):
        if i < n:
            b = i % 2
            with torch.cuda.stream(cs):
                if i >= 2: cs.wait_event(consumed[b])
                for k, v in host[i % 4].items(): dbuf[b][k].copy_(v, non_blocking=True)
                ready[b].record(cs)
        if i >= 1:
            b = (i - 1) % 2
            main.wait_event(ready[b]); out = ts.replay(buffer=b); consumed[b].record(main)
            if d2h: lh[i - 1].copy_(out["scalars"], non_blocking=True)
res["e2e loop"] = timed(e2e)
res["e2e loop, no d2h"] = timed(lambda n: e2e(n, False))
# per-step timestamps of the e2e loop: when does each copy / replay start and end on this rank?
def stamped(n=12):
    main = torch.cuda.current_stream()
    E = lambda: torch.cuda.Event(enable_timing=True)
    ready = [E(), E()]; consumed = [E(), E()]
    c0, c1, r0, r1 = [E() for _ in range(n)], [E() for _ in range(n)], [E() for _ in range(n)], [E() for _ in range(n)]
    base = E(); base.record()
    for i in range(n + 1):
        if i < n:
            b = i % 2
            with torch.cuda.stream(cs):
                if i >= 2: cs.wait_event(consumed[b])
                c0[i].record(cs)
                for k, v in host[i % 4].items(): dbuf[b][k].copy_(v, non_blocking=True)
                c1[i].record(cs); ready[b].record(cs)
        if i >= 1:
            b = (i - 1) % 2
            main.wait_event(ready[b]); r0[i - 1].record(main); ts.replay(buffer=b); r1[i - 1].record(main); consumed[b].record(main)
    torch.cuda.synchronize()
    rows = []
    for i in range(n):
        rows.append("step %2d copy %7.3f -> %7.3f   replay %7.3f -> %7.3f" % (i, base.elapsed_time(c0[i]), base.elapsed_time(c1[i]),
                                                                            base.elapsed_time(r0[i]), base.elapsed_time(r1[i])))
    return rows
if world > 1: dist.barrier()
rows = stamped()[:3]
ts.flush()
for r in range(world):
    if world > 1: dist.barrier()
    if r == rank:
        print(f"--- rank {rank}"); print("\n".join(rows), flush=True)
if rank == 0:
    print(os.environ.get("LMSFC_DP", "p2p"), {k: round(v, 4) for k, v in res.items()}, flush=True)
if world > 1:
    dist.barrier(); torch.cuda.synchronize(); ts.release(); del ts; dist.destroy_process_group()

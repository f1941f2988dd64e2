"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of `bench.py`:
picks ONE CUDA-graph replay of the train step (from the first forward kernel after an aux Adam to
the end of the next aux Adam + operand refresh) and groups its launches by kernel.
usage: python profiles/summarize_launches.py launches.csv [--detail]"""
import csv, sys, re
path = sys.argv[1]
lines = [l for l in open(path) if l.startswith('"')]
rows = [r for r in csv.DictReader(lines) if r.get("Metric Name") == "gpu__time_duration.sum"]
nm = lambda r: re.sub(r"\(.*", "", r["Kernel Name"])
us = lambda r: float(r["Metric Value"].replace(",", "")) / 1e3
prep = [i for i, r in enumerate(rows) if "adam_prep" in nm(r) or "adam_kernel" in nm(r) and False]
def start(i):
    i += 2
    while i < len(rows) and ("transpose_weight" in nm(rows[i]) or "gdn_prepare" in nm(rows[i])):
        i += 1
    return i
aux = prep[1::2]                      # every second adam_prep is the aux optimiser's
a, b = start(aux[-2]), start(aux[-1])
step = rows[a:b]
tot = sum(us(r) for r in step)
agg = {}
for r in step:
    x = agg.setdefault(nm(r)[:80], [0.0, 0]); x[0] += us(r); x[1] += 1
print(f"# one step = launches [{a},{b}) of {len(rows)}: {len(step)} launches, {tot:.1f} us (cold-cache, serialised by ncu)")
for k, (t, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{t:9.1f} us {100*t/tot:5.1f}% x{c:3d}  {k}")
if "--detail" in sys.argv:
    print()
    for r in step:
        print(f"{us(r):8.1f} {r['Grid Size']:>16} {nm(r)[:70]}")

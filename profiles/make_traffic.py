"""profiles/r2_traffic.json from an `ncu --set full` report of profiles/ncu_kernels.py:
dram__bytes_read.sum + dram__bytes_write.sum of the LAST captured launch of each roofline kernel (bench.py `kernels`).
usage: python profiles/make_traffic.py report.ncu-rep [out.json]"""
import csv, io, json, subprocess, sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
# bench.py name -> (substring of the demangled kernel name, grid x or None)
WANT = {"igemm_fwd_kernel<256>": "igemm_fwd_kernel<256", "wgrad_kernel": "wgrad_kernel", "adam_stream_kernel": "adam_stream_kernel",
        "sumsq_kernel": "sumsq_kernel", "mse_multi_kernel": "mse_multi_kernel", "gc_forward_kernel": "gc_forward_kernel",
        "gc_backward_kernel": "gc_backward_kernel", "eb_forward_kernel": "eb_forward_kernel",
        "eb_backward_kernel": "eb_backward_kernel"}
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}


def val(r, k):
    return float(r[col[k]].replace(",", "")) * UNIT.get(units[col[k]], 1)


out = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu --set full --clock-control none of "
                   "profiles/ncu_kernels.py, last captured launch of each kernel; ncu flushes the caches between its "
                   "replay passes, so this is cold-cache traffic). Summary: profiles/r2_ncu_full_kernels.txt"}
for name, sub in WANT.items():
    hits = [r for r in rows[2:] if sub in r[col["Kernel Name"]]]
    if name == "igemm_fwd_kernel<256>":          # the forward launch of the conv2 shape: 128 CTAs (dgrad launches differ)
        hits = [r for r in hits if r[col["launch__grid_size"]].replace(",", "") in ("128",)] or hits
    if not hits:
        continue
    r = hits[-1]
    out[name] = {"bytes": int(val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")),
                 "read": int(val(r, "dram__bytes_read.sum")), "write": int(val(r, "dram__bytes_write.sum")),
                 "launch_id": int(r[col["ID"]]), "duration_us": val(r, "gpu__time_duration.sum") / 1e3 if units[col["gpu__time_duration.sum"]] in ("ns", "nsecond") else r[col["gpu__time_duration.sum"]],
                 "kernel": r[col["Kernel Name"]][:60]}
dst = sys.argv[2] if len(sys.argv) > 2 else "profiles/r2_traffic.json"
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps(out, indent=1))

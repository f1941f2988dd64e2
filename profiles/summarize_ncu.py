"""Key metrics per captured launch of an `ncu --set full` report:  python profiles/summarize_ncu.py report.ncu-rep"""
import csv, subprocess, sys, io
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
KEYS = [("gpu__time_duration.sum", "duration"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs/thread"), ("launch__shared_mem_per_block_dynamic", "dyn smem"),
        ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
        ("lts__t_bytes.sum", "L2 bytes"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("smsp__cycles_active.avg", "SMSP active cycles"), ("sm__cycles_elapsed.avg.per_second", "SM clock")]
for r in rows[2:]:
    print(f"== {r[col['Kernel Name']]}  (launch id {r[col['ID']]})")
    for k, label in KEYS:
        if k in col:
            print(f"   {label:24s} {r[col[k]]:>16s} {units[col[k]]}")

#!/usr/bin/env python3
"""ncu report -> one line per profiled launch with the numbers the roofline needs.
    python tools/ncu_summary.py gpurun_out/x.ncu-rep [> profiles/rNN_x.txt]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
want = {
    "gpu__time_duration.sum": "dur_us",
    "dram__bytes_read.sum": "dram_rd",
    "dram__bytes_write.sum": "dram_wr",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pct",
    "sm__inst_executed_pipe_tensor.sum": "tensor_inst",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_pct",
    "smsp__issue_active.avg.pct": "issue_pct",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
    "l1tex__t_bytes.sum": "l1_bytes",
    "lts__t_bytes.sum": "l2_bytes",
}
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
unit_scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
print("# " + " ".join(sys.argv))
for r in rows[2:]:
    name = r[idx["Kernel Name"]][:60]
    f = {}
    for k, short in want.items():
        if k in idx and r[idx[k]] != "":
            try:
                val = float(r[idx[k]].replace(",", ""))
            except ValueError:
                continue
            f[short] = val * unit_scale.get(units[idx[k]], 1.0)
    line = f"{name:60s}"
    if "dur_us" in f:
        line += f" {f['dur_us']:10.1f} us"
    if "dram_rd" in f and "dram_wr" in f and "dur_us" in f:
        tot = f["dram_rd"] + f["dram_wr"]
        line += f"  dram {tot / 1e6:9.2f} MB ({tot / f['dur_us'] / 1e3:7.1f} GB/s)"
    for k in ("dram_pct", "tensor_pct", "issue_pct", "warps_pct", "sm_pct"):
        if k in f:
            line += f"  {k} {f[k]:5.1f}"
    for k in ("regs", "grid", "block"):
        if k in f:
            line += f"  {k} {int(f[k])}"
    print(line)

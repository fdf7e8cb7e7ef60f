#!/usr/bin/env python3
"""profiles/ncu_traffic.json from `ncu --set full` reports: DRAM bytes per launch of each workload's dominant kernel
(dram__bytes_read.sum + dram__bytes_write.sum), keyed by workload, stamped with the sha-256 of the kernel's source file
so that bench.py can tell a stale number from a current one (roofline.traffic is null when the source has changed).

    python tools/make_traffic_json.py gpurun_out/r02_tc_kernels.ncu-rep gpurun_out/r02_fh_kernels.ncu-rep"""
import csv, hashlib, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RULES = [  # (substring of the kernel name, workload, label, source file)
    ("mlp_bwd_tc_kernel", "wide4096", "mlp_bwd_tc_kernel<LOSS>", "distributed-gpu-test_b200/csrc/mlp_tc.cu"),
    ("fh_fwd_kernel", "wideFH4096", "fh_fwd_kernel", "distributed-gpu-test_b200/csrc/mlp_fh.cu"),
]


def rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units = r[0], r[1]
    idx = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for x in r[2:]:
        val = lambda k: float(x[idx[k]].replace(",", "")) * scale.get(units[idx[k]], 1)      # noqa: E731
        yield x[idx["Kernel Name"]], val("dram__bytes_read.sum") + val("dram__bytes_write.sum")


def sha(path):
    return hashlib.sha256(open(os.path.join(ROOT, path), "rb").read()).hexdigest()[:16]


out = {}
for rep in sys.argv[1:]:
    for name, b in rows(rep):
        for sub, wl, label, src in RULES:
            if sub in name:
                out[wl] = {"kernel": label, "traffic_bytes": int(b), "dtype": "bf16", "source": src, "source_sha16": sha(src),
                           "report": "profiles/" + os.path.basename(rep).replace(".ncu-rep", "").replace("r02_", "r02_ncu_") + ".txt"}
json.dump(out, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))

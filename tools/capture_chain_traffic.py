#!/usr/bin/env python
"""DRAM traffic of ONE launch of the persistent chain kernel on the headline step (Qwen3-8B, 36 layers x 4 linears,
M = 1), from an ncu capture of tools/prof_chain.py, written as the JSON bench.py reads for `roofline.traffic`:
    GIT_HASH=$(git rev-parse --short HEAD) python tools/capture_chain_traffic.py > gpurun_out/r2_chain_traffic.json
(run on the GPU box after tools/prof_chain.py has exited 0 without ncu; copy the result to profiles/)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    layers = int(os.environ.get("LAYERS", "36"))
    cmd = ["ncu", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum", "--clock-control", "none",
           "-k", "regex:chain_kernel", "-s", "1", "-c", "1", "--csv", sys.executable, os.path.join(ROOT, "tools", "prof_chain.py"),
           "--layers", str(layers), "--reps", "3"]
    out = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    lines = [l for l in out.stdout.splitlines() if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("\n".join(lines))))
    vals = {}
    for r in rows:
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(unit, 1)
        vals[r["Metric Name"]] = v * scale
    if "dram__bytes_read.sum" not in vals:
        print(out.stdout[-2000:], out.stderr[-2000:], file=sys.stderr)
        raise SystemExit("ncu produced no dram metrics")
    sys.path.insert(0, ROOT)
    import bench
    algo = layers * sum(bench.algo_bytes(1, K, N) for _, K, N in bench.QWEN3_8B)
    total = vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"]
    print(json.dumps({"kernel": "chain_kernel<1,1>", "git": os.environ.get("GIT_HASH"), "layers": layers,
                      "what": "dram__bytes_read.sum + dram__bytes_write.sum of one launch (ncu, cold serialised replay)",
                      "dram_read": int(vals["dram__bytes_read.sum"]), "dram_write": int(vals["dram__bytes_write.sum"]),
                      "mean_per_launch": int(total), "algorithmic_bytes_per_launch": int(algo),
                      "ratio": round(total / algo, 4), "ncu_duration_us": round(vals.get("gpu__time_duration.sum", 0), 1)}))


if __name__ == "__main__":
    main()

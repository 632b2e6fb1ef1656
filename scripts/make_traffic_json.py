#!/usr/bin/env python
"""profiles/r02_traffic.json from an `ncu --set full` report of one frame-step (scripts/gpu/r2_ncu_chain.sh):
dram__bytes_read.sum + dram__bytes_write.sum per launch of the two persistent kernels bench.py reports a roofline for,
stamped with the hash of the kernel sources (bench.py prints `traffic: null` when the sources have changed since).

    python scripts/make_traffic_json.py gpurun_out/r2_prof_chain.ncu-rep [--batch 64]
"""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_traffic.json"))
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        return float(r[ix[name]].replace(",", "")) * UNIT[units[ix[name]]]

    launches = []
    for r in data:
        if "k_chain" not in r[ix["Kernel Name"]]:
            continue
        launches.append({"dram_bytes": int(val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")),
                         "dram_bytes_read": int(val(r, "dram__bytes_read.sum")),
                         "duration_us": float(r[ix["gpu__time_duration.sum"]].replace(",", "")) *
                         {"us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}[units[ix["gpu__time_duration.sum"]]]})
    if len(launches) < 2:
        raise SystemExit("expected the head's and the decoder stage's k_chain launches in the report")
    launches.sort(key=lambda e: -e["duration_us"])          # the head's solver loop is the long one
    import bench
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    out = {"source_sha16": bench.kernel_source_hash(), "sources": bench.TRAFFIC_SOURCES, "captured_after_commit": commit,
           "ddpm_steps": bench.DDPM_STEPS, "report": os.path.basename(args.report),
           "how": "ncu --set full --clock-control none, one frame-step after warm-up (scripts/gpu/r2_ncu_chain.sh)",
           "k_chain_head": dict(launches[0], batch=args.batch),
           "k_chain_decoder_stage0": dict(launches[1], batch=args.batch)}
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

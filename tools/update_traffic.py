#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --set full` capture: mean dram__bytes_read.sum + dram__bytes_write.sum per launch of each kernel
named on the command line, stamped with the hash of the CUDA sources the library was built from (engine.kernel_source_hash()). bench.py
refuses a stamp that does not match the sources it runs (roofline.traffic = null), so a capture can never outlive its kernel.

  python tools/update_traffic.py gpurun_out/extend.ncu-rep k_extend [k_intersect ...] [--note "..."]"""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("kernels", nargs="+")
    ap.add_argument("--note", default="")
    ap.add_argument("--hash", default=None, help="source hash the capture was taken on (default: the sources as they are now)")
    args = ap.parse_args()
    from vermilion_b200 import engine
    out = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, data = rows[0], rows[2:]
    ir, iw, it, ik = (hdr.index(k) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "Kernel Name"))
    units = rows[1]

    def to_bytes(v, u):
        return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]

    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        tr = json.load(open(path))
    except Exception:
        tr = {}
    for k in args.kernels:
        sel = [r for r in data if k in r[ik]]
        if not sel:
            print(f"no launch of {k} in {args.report}", file=sys.stderr)
            continue
        tot = [to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw]) for r in sel]
        tr[k] = {"dram_bytes_per_launch": int(sum(tot) / len(tot)), "launches": len(sel), "source_hash": args.hash or engine.kernel_source_hash(),
                 "source": f"{os.path.basename(args.report)}: mean over {len(sel)} launches of dram__bytes_read.sum + dram__bytes_write.sum "
                           f"(ncu --set full --clock-control none). {args.note}".strip()}
        print(k, tr[k])
    json.dump(tr, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()

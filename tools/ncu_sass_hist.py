"""Dynamic SASS opcode histogram (instructions executed and stall samples per opcode) from an ncu report:
   python tools/ncu_sass_hist.py gpurun_out/x.ncu-rep [units]   (units: divide counts by this many, e.g. slot-warps)"""
import csv
import subprocess
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = None
    inst, samp = defaultdict(float), defaultdict(float)
    for r in rows:
        if r and r[0] == "Address":
            hdr = r
            continue
        if not hdr or len(r) < len(hdr) or not r[0].startswith("0x"):
            continue
        d = dict(zip(hdr, r))
        toks = d["Source"].split()
        op = toks[0]
        if op.startswith("@"):
            op = "@" + toks[1]
        op = op.split(".")[0] + ("." + ".".join(toks[1 if not toks[0].startswith("@") else 1].split(".")[1:2]) if False else "")
        inst[op] += float(d["Instructions Executed"] or 0)
        samp[op] += float(d["# Samples"] or 0)
    ti, ts = sum(inst.values()), sum(samp.values())
    print(f"total instructions {ti:.0f} ({ti / units:.2f} per unit), samples {ts:.0f}")
    for op, n in sorted(inst.items(), key=lambda x: -x[1])[:28]:
        print(f"{op:14s} {n:14.0f} {n / units:8.2f}/unit {100 * n / ti:5.1f}% inst {100 * samp[op] / max(ts, 1):5.1f}% samples")


if __name__ == "__main__":
    main()

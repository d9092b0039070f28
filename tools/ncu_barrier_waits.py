"""Where a barrier-bound kernel waits: per BAR.SYNC of an ncu report (captured with --import-source on) the warp-stall samples that
piled up at the barrier (attributed by ncu to the instructions right after it), with the source line of the barrier from nvdisasm.
   python tools/ncu_barrier_waits.py report.ncu-rep object.o kernel_substring [top]"""
import csv
import re
import subprocess
import sys
import tempfile
import os


def main():
    rep, obj, kname = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ia, isrc, isamp = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples")
    ins = [(int(r[ia], 16), r[isrc].strip(), int(r[isamp] or 0)) for r in rows[2:] if len(r) > isamp and r[ia].startswith("0x")]
    base = ins[0][0]
    # source lines per offset from nvdisasm
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cubin)], capture_output=True, text=True).stdout
    line_of, cur, on = {}, None, False
    for l in dis.split("\n"):
        if l.startswith(".text.") or "Function :" in l or l.startswith("\t.section"):
            on = kname in l if (".text." in l) else on
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
        if m and on and cur:
            line_of[int(m.group(1), 16)] = cur
    total = sum(s for _, _, s in ins)
    waits = []
    for i, (a, src, s) in enumerate(ins):
        if "BAR.SYNC" in src:
            w = sum(x[2] for x in ins[i:i + 3])     # samples at the barrier and the two instructions after it
            waits.append((w, a - base, line_of.get(a - base)))
    print(f"total samples {total}; barriers {len(waits)}; samples at barriers {sum(w for w, _, _ in waits)} ({100 * sum(w for w, _, _ in waits) / total:.1f} %)")
    for w, off, ln in sorted(waits, reverse=True)[:top]:
        print(f"{100 * w / total:5.1f}%  +0x{off:05x}  {ln}")


if __name__ == "__main__":
    main()

"""Top source lines by warp-stall samples from an ncu report captured with --import-source on (kernel built with -lineinfo):
   python tools/ncu_hot_lines.py gpurun_out/x.ncu-rep [top_n]"""
import csv
import subprocess
import sys


def main():
    path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows, fname, hdr = [], "", None
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] in ("File Name", "File Path"):
            fname = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit():
            d = dict(zip(hdr, r))
            try:
                n = int(d["# Samples"])
            except Exception:
                continue
            if n:
                stalls = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)}
                rows.append((n, fname, int(r[0]), r[1].strip()[:110], int(d["Instructions Executed"] or 0), stalls))
    tot = sum(r[0] for r in rows)
    print(f"total samples {tot}")
    for n, f, ln, src, inst, st in sorted(rows, key=lambda r: r[:3], reverse=True)[:top]:
        tops = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda x: -x[1])[:3])
        print(f"{100 * n / tot:5.1f}%  {f}:{ln:<4d} inst {inst:>10d}  [{tops}]  {src}")


if __name__ == "__main__":
    main()

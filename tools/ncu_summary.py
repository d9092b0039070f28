"""Print the metrics that matter from an ncu report: python tools/ncu_summary.py gpurun_out/x.ncu-rep [more-metric-substrings...]"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__waves_per_multiprocessor", "sm__inst_executed_pipe_fma",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.max", "sm__pipe_tensor", "sm__inst_executed_pipe_tensor", "smsp__average_warp", "smsp__warp_issue_stalled",
        "sm__pipe_fp64", "launch__occupancy", "smsp__cycles_active.avg", "sm__cycles_active.avg"]


def main():
    path = sys.argv[1]
    extra = sys.argv[2:]
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[hdr.index("Kernel Name")][:100] if "Kernel Name" in hdr else "")
        for i, k in enumerate(hdr):
            if any(k.startswith(x) for x in KEYS) or any(x in k for x in extra):
                if "stalled" in k and not k.endswith("_warps_active.ratio") and "pct" not in k:
                    continue
                print(f"  {k:100s} {r[i]:>18s} {units[i]}")


if __name__ == "__main__":
    main()

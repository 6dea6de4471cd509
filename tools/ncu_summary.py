"""Per-launch summary of an ncu --set full report (read on the CPU box):  python tools/ncu_summary.py file.ncu-rep"""
import csv
import subprocess
import sys

WANT = ["Kernel Name", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum", "sm__cycles_elapsed.max"]


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    idx = [(w, h.index(w)) for w in WANT if w in h]
    for r in rows[2:]:
        print("---")
        for w, i in idx:
            print(f"  {w:70s} {r[i][:90]} {units[i]}")


if __name__ == "__main__":
    main()

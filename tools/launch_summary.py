"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total time and share.
    python tools/launch_summary.py gpurun_out/launches.csv [first_id last_id]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 60
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 14 and r[0].isdigit()]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r[12] != "gpu__time_duration.sum" or not (lo <= int(r[0]) < hi):
            continue
        name = re.sub(r"\((?!.*<).*", "", r[4]).replace("unnamed>::", "").replace("void ", "")
        name = re.sub(r"\(.*", "", name)
        agg[name][0] += 1
        agg[name][1] += float(r[14])
    tot = sum(v[1] for v in agg.values())
    n = sum(v[0] for v in agg.values())
    print(f"{n} launches, {tot / 1e6:.3f} ms (serialised, cold-cache: compare shares)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1] / 1e6:9.3f} ms {100 * v[1] / tot:5.1f}%  n={v[0]:4d}  avg {v[1] / v[0] / 1e3:8.1f} us  {k}")


if __name__ == "__main__":
    main()

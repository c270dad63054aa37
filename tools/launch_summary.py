"""Summarise an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`): per-kernel launches, total time, share.
  python tools/launch_summary.py gpurun_out/launches.csv [--from-first KERNEL_SUBSTRING]"""
import collections
import csv
import re
import sys


def rows(path):
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.reader(lines):
        if r[0] == "ID" or len(r) < 15:
            continue
        name = re.sub(r"\(.*", "", r[4]).replace("void ", "")
        yield name, r[7], r[8], float(r[14].replace(",", "")) / 1e3


def main():
    path = sys.argv[1]
    start = sys.argv[3] if len(sys.argv) > 3 and sys.argv[2] == "--from-first" else None
    data = list(rows(path))
    if start:
        idx = next(i for i, r in enumerate(data) if start in r[0])
        data = data[idx:]
    agg = collections.OrderedDict()
    for name, _, _, us in data:
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1; a[1] += us
    tot = sum(a[1] for a in agg.values())
    print("%d launches, %.2f ms of kernel time" % (len(data), tot / 1e3))
    print("%-34s %9s %10s %7s %9s" % ("kernel", "launches", "total_ms", "share", "avg_us"))
    for name, (cnt, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-34s %9d %10.2f %6.1f%% %9.1f" % (name, cnt, us / 1e3, 100 * us / tot, us / cnt))


if __name__ == "__main__":
    main()

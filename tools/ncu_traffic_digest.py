"""Digest of an ncu raw-page CSV taken with tools/prof_sort.py: DRAM traffic of the NTT launches of one sort against their
algorithmic bytes.  Every ntt_kernel CTA owns one tile of 2^T words (T = min(log2 of the transform length, 10)); a pass of
a two-pass transform reads and writes each word once, and the algorithmic bytes of a transform (read once, write once)
belong to the PAIR of passes, so one pass launch is credited grid x 2^T x 8 bytes (a single-pass transform 16).
  python tools/ncu_traffic_digest.py gpurun_out/ntt_in_sort.csv > profiles/r02_ntt_traffic_in_sort.json"""
import csv, json, re, sys

def rows(path):
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    r = csv.reader(lines)
    head = next(r)
    units = next(r)
    for row in r:
        if len(row) == len(head):
            yield dict(zip(head, row)), dict(zip(head, units))

def num(s):
    return float(s.replace(",", "")) if s not in ("", "n/a") else 0.0

def scale(unit):
    return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1.0)

def main():
    per = {}
    for row, units in rows(sys.argv[1]):
        name = row["Kernel Name"]
        m = re.search(r"ntt_kernel<\(?(?:bool\))?(\w+), *\(?(?:int\))?(\d+), *\(?(?:int\))?(\d+)(?:, *\(?(?:int\))?\d+)?>", name)   # the 4th argument (row kinds / fused work) is folded
        if not m:
            continue
        key = "ntt_kernel<%s,%s,%s>" % m.groups()
        grid = int(row["Grid Size"].strip("()").split(",")[0])
        logn = int(m.group(2))
        single = logn != 0 and logn <= 10
        T = min(logn, 10) if logn else 10
        alg = grid * (1 << T) * (16 if single else 8)
        rd = num(row["dram__bytes_read.sum"]) * scale(units["dram__bytes_read.sum"])
        wr = num(row["dram__bytes_write.sum"]) * scale(units["dram__bytes_write.sum"])
        us = num(row["gpu__time_duration.sum"]) * scale(units["gpu__time_duration.sum"])
        d = per.setdefault(key, {"launches": 0, "alg_bytes": 0.0, "dram_read": 0.0, "dram_write": 0.0, "us": 0.0})
        d["launches"] += 1; d["alg_bytes"] += alg; d["dram_read"] += rd; d["dram_write"] += wr; d["us"] += us
    tot = {k: sum(d[k] for d in per.values()) for k in ("launches", "alg_bytes", "dram_read", "dram_write", "us")}
    out = {"source": "ncu --profile-from-start off --clock-control none -k regex:ntt_kernel --metrics dram__bytes_read.sum,dram__bytes_write.sum,"
                     "gpu__time_duration.sum, python tools/prof_sort.py: every NTT launch of ONE steady-state DirectSort N=128 sort "
                     "(third sort of the process: plaintext cache and memory pool warm)",
           "per_kernel": per, "total": tot,
           "traffic_over_algorithmic": (tot["dram_read"] + tot["dram_write"]) / tot["alg_bytes"] if tot["alg_bytes"] else None,
           "avg_alg_bytes_per_launch": tot["alg_bytes"] / tot["launches"] if tot["launches"] else None,
           "avg_dram_bytes_per_launch": (tot["dram_read"] + tot["dram_write"]) / tot["launches"] if tot["launches"] else None,
           "note": "per-launch times under ncu are serialised (no overlap with the plaintext side stream); the byte counters are what this file is for"}
    print(json.dumps(out, indent=1))

main()

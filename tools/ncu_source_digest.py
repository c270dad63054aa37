"""Per-opcode and per-instruction stall digest of an `ncu --page source --csv --print-source sass` export.
  python tools/ncu_source_digest.py src.csv [kernel_index ...]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = [int(a) for a in sys.argv[2:]]
ks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; ks.append(cur); continue
    if r and r[0] == "Address":
        cur["hdr"] = r; continue
    if cur is not None and len(r) > 10:
        cur["rows"].append(r)
for n, k in enumerate(ks):
    if want and n not in want:
        continue
    h = k["hdr"]; si = h.index("# Samples"); ii = h.index("Instructions Executed")
    tot = sum(int(r[si]) for r in k["rows"])
    ex = [r for r in k["rows"] if int(r[ii]) > 0]
    print("[%d] %s: samples %d, executed SASS lines %d, warp instructions %d" % (n, k["name"][:70], tot, len(ex), sum(int(r[ii]) for r in ex)))
    byop, sam = collections.Counter(), collections.Counter()
    for r in ex:
        f = r[1].split()
        op = (f[1] if f[0].startswith("@") else f[0]).split(".")[0]
        byop[op] += int(r[ii]); sam[op] += int(r[si])
    for op, c in byop.most_common(24):
        print("   %-8s inst %9d  samples %6d (%.1f%%)" % (op, c, sam[op], 100 * sam[op] / max(tot, 1)))
    stall_cols = [i for i, x in enumerate(h) if x.startswith("stall_") and "(Not" not in x]
    agg = collections.Counter()
    for r in ex:
        for i in stall_cols:
            agg[h[i]] += int(r[i])
    print("   stall totals:", ", ".join("%s %d" % kv for kv in agg.most_common(9)))
    for r in sorted(ex, key=lambda r: -int(r[si]))[:28]:
        st = sorted([(int(r[i]), h[i][6:]) for i in stall_cols], reverse=True)[:2]
        print("   ", r[si].rjust(5), r[1].strip()[:64].ljust(64), st)

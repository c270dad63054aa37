"""Instruction-mix digest of cuobjdump -sass output, per kernel: total instructions and the counts of the mnemonics that
show which pipe a kernel lives on (DFMA/DMUL/DADD = FP64, IMAD/IADD3 = integer, LDG/STG widths, BAR, griddepcontrol).
  cuobjdump -sass sorting-fhe_b200/csrc/build/ntt.o | python tools/sass_histogram.py ntt_kernel > profiles/r02_sass_ntt.txt"""
import collections, re, sys
pat = sys.argv[1] if len(sys.argv) > 1 else ""
cur, hist = None, {}
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); hist[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
keys = ["DFMA", "DMUL", "DADD", "IMAD", "IADD3", "LOP3", "SHF", "LDG", "STG", "LDS", "STS", "BAR", "CCTL", "ACQBULK", "DEPBAR"]
for fn, h in hist.items():
    if pat and pat not in fn:
        continue
    tot = sum(h.values())
    fam = collections.Counter()
    for op, c in h.items():
        fam[op.split(".")[0]] += c
    wide = {op: c for op, c in h.items() if op.startswith(("LDG", "STG", "LDS", "STS")) }
    print(fn)
    print("  total %d;  " % tot + "  ".join("%s %d" % (k, fam[k]) for k in keys if fam[k]))
    print("  memory ops: " + ", ".join("%s x%d" % kv for kv in sorted(wide.items())))

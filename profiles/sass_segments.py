"""Split the ncu source page (SASS) of a kernel into segments between barriers and report, per segment,
executed warp instructions, stall samples and the opcode mix.  Usage:
    ncu -i X.ncu-rep --page source --csv > /tmp/src.csv ; python profiles/sass_segments.py /tmp/src.csv
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
I = {n: i for i, n in enumerate(hdr)}
seg, segs = [], []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    seg.append(r)
    op = r[I["Source"]].split()
    if any(o.startswith(("BAR", "SYNCS", "WARPSYNC.ALL")) for o in op[:2]) and "BAR" in r[I["Source"]]:
        segs.append(seg)
        seg = []
segs.append(seg)
tot_i = sum(int(r[I["Instructions Executed"]]) for s in segs for r in s)
tot_s = sum(int(r[I["# Samples"]]) for s in segs for r in s)
print(f"total warp instructions {tot_i}, samples {tot_s}, segments {len(segs)}")
print(f"{'seg':>3} {'sass':>5} {'inst':>12} {'inst%':>6} {'samp%':>6} {'bar%':>6} {'ssb%':>6} {'lsb%':>6}  top opcodes")
for k, s in enumerate(segs):
    n = sum(int(r[I["Instructions Executed"]]) for r in s)
    sm = sum(int(r[I["# Samples"]]) for r in s)
    sb = sum(int(r[I["stall_barrier"]]) for r in s)
    ss = sum(int(r[I["stall_short_sb"]]) for r in s)
    ls = sum(int(r[I["stall_long_sb"]]) for r in s)
    ops = collections.Counter()
    for r in s:
        toks = r[I["Source"]].split()
        op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
        ops[op.split(".")[0]] += int(r[I["Instructions Executed"]])
    top = " ".join(f"{o}:{100 * c // max(n, 1)}%" for o, c in ops.most_common(6))
    print(f"{k:3d} {len(s):5d} {n:12d} {100 * n / tot_i:6.2f} {100 * sm / max(tot_s, 1):6.2f} {100 * sb / max(sm, 1):6.1f} {100 * ss / max(sm, 1):6.1f} {100 * ls / max(sm, 1):6.1f}  {top}")

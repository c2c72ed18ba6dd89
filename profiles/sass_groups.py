"""Group the ncu source page (SASS) of a kernel into runs of instructions with the same execution count and print,
per run, instructions per warp-round, stall samples and the opcode mix: attributes instruction counts to loops.
Usage: ncu -i X.ncu-rep --page source --csv > /tmp/src.csv ; python profiles/sass_groups.py /tmp/src.csv <rounds*warps> [min]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
base = float(sys.argv[2])
floor = float(sys.argv[3]) if len(sys.argv) > 3 else 8.0
hdr = rows[1]
I = {n: i for i, n in enumerate(hdr)}
cur, grp, out = None, [], []
for idx, r in enumerate(rows[2:]):
    if len(r) < len(hdr):
        continue
    n = int(r[I["Instructions Executed"]])
    toks = r[I["Source"]].split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    key = round(n / base, 2)
    if cur is None or abs(key - cur) > 0.02 * max(cur, 0.05):
        if grp:
            out.append((cur, grp))
        grp, cur = [], key
    grp.append((idx, op, n, int(r[I["# Samples"]])))
out.append((cur, grp))
for key, g in out:
    s = sum(x[2] for x in g) / base
    if s >= floor:
        ops = collections.Counter(x[1].split(".")[0] for x in g)
        print(f"sass {g[0][0]:5d}-{g[-1][0]:5d} n={len(g):4d} exec={key:6.2f} instr/warp-round={s:7.1f} samples={sum(x[3] for x in g):6d}  "
              + " ".join(f"{o}:{c}" for o, c in ops.most_common(7)))

"""Per-SOURCE-LINE summary of an ncu report taken with --import-source on (kernel compiled with -lineinfo).

    python profiles/src_lines.py <report.ncu-rep> <warp-rounds in the launch> [top]

Prints, for the source lines with the most executed warp instructions: instructions per warp-round, share of all
instructions, share of stall samples and the dominant stall reasons.  (warp-rounds = rounds x warps per round.)"""
import collections
import csv
import subprocess
import sys

rep, wr = sys.argv[1], float(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fpath = None
hdr = None
agg = collections.OrderedDict()
stall_names = None
for row in csv.reader(out.splitlines()):
    if not row:
        continue
    if row[0] == "File Path":
        fpath = row[1].split("/")[-1]
        continue
    if row[0] == "Line No":
        hdr = row
        I = {}
        for i, n in enumerate(hdr):
            I.setdefault(n, i)
        stall_names = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
        continue
    if hdr is None or len(row) < len(hdr) or row[0] == "":
        continue
    try:
        line = int(row[0])
    except ValueError:
        continue
    a = agg.setdefault((fpath, line), {"src": row[1].strip()[:90], "inst": 0, "samp": 0, "st": collections.Counter()})
    num = lambda v: int(v) if v.lstrip("-").isdigit() else 0      # noqa: E731
    a["inst"] += num(row[I["Instructions Executed"]])
    a["samp"] += num(row[I["# Samples"]])
    for n in stall_names:
        a["st"][n[6:]] += num(row[I[n]])
ti = sum(a["inst"] for a in agg.values())
ts = sum(a["samp"] for a in agg.values())
print(f"# {rep}: {ti} warp instructions = {ti / wr:.0f} per warp-round, {ts} stall samples")
print(f"{'file:line':32s} {'inst/wr':>8} {'inst%':>6} {'samp%':>6}  top stalls | source")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["inst"])[:top]:
    st = " ".join(f"{n}:{100 * c // max(a['samp'], 1)}%" for n, c in a["st"].most_common(3) if c)
    print(f"{k[0] + ':' + str(k[1]):32s} {a['inst'] / wr:8.1f} {100 * a['inst'] / ti:6.2f} {100 * a['samp'] / max(ts, 1):6.2f}  {st} | {a['src']}")
print("# lines by stall samples")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["samp"])[:20]:
    st = " ".join(f"{n}:{100 * c // max(a['samp'], 1)}%" for n, c in a["st"].most_common(4) if c)
    print(f"{k[0] + ':' + str(k[1]):32s} {a['inst'] / wr:8.1f} {100 * a['inst'] / ti:6.2f} {100 * a['samp'] / max(ts, 1):6.2f}  {st} | {a['src']}")

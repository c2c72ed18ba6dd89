"""Dynamic opcode histogram of an ncu report taken with --set full (SASS view of the source page).

    python profiles/opcode_hist.py <report.ncu-rep> <warp-rounds in the launch> [top]

Prints executed warp instructions per opcode (modifiers after the first dot kept for the top entries), per warp-round
and as a share: what the issue slots of an instruction-bound kernel are spent on."""
import collections
import csv
import subprocess
import sys

rep, wr = sys.argv[1], float(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
hdr = None
full = collections.Counter()
base = collections.Counter()
for row in csv.reader(out.splitlines()):
    if not row:
        continue
    if "Instructions Executed" in row and "Source" in row:
        hdr = row
        isrc, iex = row.index("Source"), row.index("Instructions Executed")
        continue
    if hdr is None or len(row) < len(hdr):
        continue
    try:
        n = int(row[iex])
    except ValueError:
        continue
    toks = row[isrc].replace("{", " ").split()
    if not toks:
        continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    op = op.rstrip(";")
    full[op] += n
    base[op.split(".")[0]] += n
if hdr is None:
    print("no SASS table found; first lines of the ncu output:")
    print("\n".join(out.splitlines()[:15]))
    sys.exit(1)
total = sum(base.values())
print(f"# {rep}: {total} warp instructions = {total / wr:.0f} per warp-round")
print("# by opcode")
for op, n in base.most_common(top):
    print(f"{op:24s} {n / wr:9.1f} {100 * n / total:6.2f} %")
print("# by opcode with modifiers")
for op, n in full.most_common(top):
    print(f"{op:32s} {n / wr:9.1f} {100 * n / total:6.2f} %")

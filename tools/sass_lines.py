#!/usr/bin/env python
"""Per-source-line SASS statistics of one kernel (static counts; needs -lineinfo objects).

    python tools/sass_lines.py <object.o> <substring of the mangled kernel name> [top]

Prints the instruction count, spill stores / loads (STL / LDL) per source line and the opcode histogram:
what to look at before spending GPU time on a kernel (register spills inside hot loops, code size)."""
import collections
import os
import re
import subprocess
import sys
import tempfile


def main():
    obj, pat = os.path.abspath(sys.argv[1]), sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    with tempfile.TemporaryDirectory() as d:
        subprocess.check_call(["cuobjdump", "-xelf", "all", obj], cwd=d, stdout=subprocess.DEVNULL)
        cubins = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")]
        text = "".join(subprocess.run(["nvdisasm", "-g", "-c", c], capture_output=True, text=True).stdout for c in cubins)
    lines = text.split("\n")
    start = end = None
    for i, l in enumerate(lines):
        if l.startswith("\t.section") and ".text." in l:
            if start is None and pat in l:
                start = i
            elif start is not None:
                end = i
                break
    sec = lines[start:end]
    cur = None
    tot, stl, ldl, ops = (collections.Counter() for _ in range(4))
    n = 0
    for l in sec:
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?(\S+)", l)
        if m:
            n += 1
            op = m.group(1).rstrip(";")
            tot[cur] += 1
            ops[op.split(".")[0]] += 1
            if op.startswith("STL"):
                stl[cur] += 1
            if op.startswith("LDL"):
                ldl[cur] += 1
    print(f"{n} instructions ({16 * n / 1024:.0f} KB)")
    print("spill stores by line:", sorted(stl.items(), key=lambda kv: str(kv[0])))
    print("spill loads by line: ", sorted(ldl.items(), key=lambda kv: str(kv[0])))
    print("opcodes:", ", ".join(f"{k}:{v}" for k, v in ops.most_common(30)))
    for k, v in tot.most_common(top):
        print(f"{str(k):50s} {v}")


if __name__ == "__main__":
    main()

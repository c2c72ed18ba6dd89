"""Update profiles/traffic.json from an ncu summary written by profiles/summarize.py.

    python profiles/make_traffic.py <variant> <rounds in the captured launch> <profiles/xxx_full.txt> <bytes per round> <commit>
"""
import json
import os
import re
import sys

variant, rounds, path, algo, commit = sys.argv[1], int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5]
text = open(path).read()


def val(name):
    m = re.search(name + r"\s+([\d.]+)\s+(\w+)", text)
    return float(m.group(1)) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[m.group(2)]


total = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json")
d = json.load(open(out))
d[variant] = {"rounds": rounds, "dram_bytes": total, "dram_bytes_per_round": total / rounds,
              "algorithmic_bytes_per_round": algo, "source": path, "commit": commit}
json.dump(d, open(out, "w"), indent=2)
print(variant, total / rounds, "bytes per round,", total / rounds / algo, "x algorithmic")

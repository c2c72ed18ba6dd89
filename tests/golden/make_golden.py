"""Generate tests/golden/notebook_golden.npz by running the REFERENCE's own code.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):   python tests/golden/make_golden.py

The float form of the consensus lives in the notebook
  /root/reference/contract/drafts/beta_kumaraswamy_algorithm_demo copy.ipynb
Its definition cells (3, 4, 12, 13, 17, 18) need only numpy + itertools; cell 1
imports seaborn/matplotlib, which are absent, so it is skipped and the three
names it would provide are injected.  Nothing is copied from the notebook: the
cells are exec'd from where they lie and only their numeric outputs are stored.
"""
import io
import json
import contextlib
import os
import re
import sys
from itertools import compress
from typing import List, Tuple

import numpy as np

REF = "/root/reference"
NB = os.path.join(REF, "contract/drafts/beta_kumaraswamy_algorithm_demo copy.ipynb")
HERE = os.path.dirname(os.path.abspath(__file__))


def load_notebook_namespace():
    nb = json.load(open(NB))
    ns = {"np": np, "List": List, "Tuple": Tuple, "compress": compress}
    for idx in (3, 4, 12, 13, 17, 18):
        exec("".join(nb["cells"][idx]["source"]), ns)
    return ns, nb


def main():
    ns, nb = load_notebook_namespace()
    out = {}

    # ---- known answers printed in the notebook (cells 14, 15)
    txt14 = "".join(nb["cells"][14]["outputs"][0]["text"])
    kat1d = np.array(json.loads(txt14.splitlines()[1]), dtype=np.int64)
    txt15 = "".join(nb["cells"][15]["outputs"][0]["text"])
    kat2d = np.array([[int(a), int(b)] for a, b in re.findall(r"array!\[(\d+), (\d+)\]", txt15)],
                     dtype=np.int64)
    out["kat1d_wsad"] = kat1d
    out["kat2d_wsad"] = kat2d
    out["kat1d_printed"] = np.array(float(re.search(r": ([\d.]+)", txt14).group(1)))
    out["kat2d_printed"] = np.array(float(re.search(r": ([\d.]+)", txt15).group(1)))
    out["kat1d_expected_reliability"] = np.array(ns["expected_reliability"](kat1d * 1e-6))
    out["kat2d_expected_reliability"] = np.array(ns["expected_reliability"](kat2d * 1e-6))

    # ---- seeded rounds through the reference's estimator
    shapes = [(7, 1, 2, 64), (7, 2, 2, 64), (7, 6, 2, 32), (20, 1, 2, 32), (20, 2, 4, 32),
              (20, 1, 15, 16), (32, 8, 6, 16), (64, 16, 12, 8), (256, 128, 51, 2)]
    rng = np.random.RandomState(20240607)
    meta = []
    for (n, d, f, t) in shapes:
        key = f"n{n}_d{d}_f{f}"
        preds = np.empty((t, n, d), dtype=np.float64)
        masks = np.empty((t, n), dtype=bool)
        cons = np.empty((t, d), dtype=np.float64)
        ranks = np.empty((t, n), dtype=np.int64)
        nranks = np.empty((t, n), dtype=np.float64)
        erel = np.empty((t,), dtype=np.float64)
        for k in range(t):
            np.random.seed(rng.randint(0, 2**31 - 1))
            if d == 1:
                vals, _truth = ns["generate_beta_oracles"](n, f, 10, 10)
                p = np.array(vals, dtype=np.float64)[:, None]
                arg = list(vals)                      # the 1-D path takes a list of floats
            elif d == 2:
                p, _truth = ns["generate_2d_beta_oracles"](n, f, [10, 10], [15, 30])
                arg = p
            else:
                honest = np.random.beta(10, 15, size=(n, d))
                bad = np.random.uniform(0, 1, size=(n, d))
                isbad = np.zeros(n, dtype=bool)
                isbad[np.random.permutation(n)[:f]] = True
                p = np.where(isbad[:, None], bad, honest)
                # store as float32-representable values so the fp32 engine sees identical inputs
                p = p.astype(np.float32).astype(np.float64)
                arg = p
            m = ns["identify_failing_oracles"](arg, f)
            c = ns["restricted_median"](arg, m)
            dev = [np.linalg.norm(np.asarray(x) - np.median(arg, axis=0)) for x in arg]
            nr, r = ns["rank_array"](dev)
            preds[k] = p
            masks[k] = np.array(m, dtype=bool)
            cons[k] = np.atleast_1d(c)
            ranks[k] = r
            nranks[k] = nr
            erel[k] = ns["expected_reliability"](np.asarray(arg))
        out[key + "_pred"] = preds
        out[key + "_mask"] = masks
        out[key + "_consensus"] = cons
        out[key + "_ranks"] = ranks
        out[key + "_nranks"] = nranks
        out[key + "_expected_reliability"] = erel
        meta.append((n, d, f, t))
    out["shapes"] = np.array(meta, dtype=np.int64)

    # ---- the reference's Monte-Carlo benchmark at large K (statistical anchor for the
    #      on-device generator + engine; the published K=300 table is in BASELINE.md)
    stats = []
    for (n, f) in [(7, 2), (20, 2), (20, 15), (7, 0), (7, 1)]:      # F = 0, 1: anchors of the failing-fraction sweep (config 2)
        for a in (10, 20, 30, 100):
            np.random.seed(1000 * n + 10 * f + a)
            buf = io.StringIO()
            k = 20000 if n == 7 else 8000
            with contextlib.redirect_stdout(buf):
                ns["benchmark"](a, (a, a), n, f, k)
            line = buf.getvalue()
            succ, rel = re.findall(r": ?([\d.]+) %", line)
            stats.append((n, f, a, k, float(succ), float(rel)))
            print(line.strip(), file=sys.stderr)
    out["benchmark_stats"] = np.array(stats, dtype=np.float64)  # n, f, a, K, success %, reliability %

    np.savez_compressed(os.path.join(HERE, "notebook_golden.npz"), **out)
    print("wrote notebook_golden.npz with", len(out), "arrays", file=sys.stderr)


if __name__ == "__main__":
    main()

"""Build an A/B variant of libsvoc.so: recompile the listed translation units with extra -D flags and link
them with the objects of the regular build.  The result goes to <package>/build/ab/libsvoc_<tag>.so (git-ignored,
travels to the GPU box) and is selected at run time with SVOC_LIB=<path>.

    python tools/ab_build.py <tag> [-DNAME=VALUE ...] [--tu consensus_f32.cu ...]
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "stochastic-vector-oracle-consensus_b200")
sys.path.insert(0, PKG)
import build as B  # noqa: E402


def main():
    tag = sys.argv[1]
    defs = [a for a in sys.argv[2:] if a.startswith("-D")]
    tus = []
    if "--tu" in sys.argv:
        tus = sys.argv[sys.argv.index("--tu") + 1:]
    tus = tus or ["consensus_f32.cu"]
    B.build()
    out_dir = os.path.join(B.OBJ, "ab", tag)
    os.makedirs(out_dir, exist_ok=True)
    objs = []
    procs = []
    for name in sorted(os.listdir(B.CSRC)):
        if not name.endswith(".cu"):
            continue
        if name in tus:
            obj = os.path.join(out_dir, name[:-3] + ".o")
            procs.append(subprocess.Popen([B.NVCC] + B.FLAGS + defs + ["-c", os.path.join(B.CSRC, name), "-o", obj]))
        else:
            obj = os.path.join(B.OBJ, name[:-3] + ".o")
        objs.append(obj)
    for p in procs:
        if p.wait() != 0:
            raise SystemExit("nvcc failed")
    out = os.path.join(B.OBJ, "ab", f"libsvoc_{tag}.so")
    subprocess.check_call([B.NVCC, "-shared", "-o", out] + objs + ["-lcudart"])
    print(out)


if __name__ == "__main__":
    main()

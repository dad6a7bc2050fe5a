"""Static SASS mnemonic counts per kernel of the built library (cuobjdump -sass): the evidence that the hot
kernels use tcgen05 / TMA / TMEM (profiles/r02_sass_hist.txt).  Usage: python tools/sass_hist.py > profiles/..."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sigsvgd_b200", "libsigsvgd_b200.so")
COLS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "FFMA", "FFMA2", "DFMA", "MUFU", "F2FP",
        "LDS", "STS", "LDG", "STG", "REDG", "ATOMS", "ATOMG", "SHFL", "BAR"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(name):
    name = re.sub(r"\(anonymous namespace\)::|sigsvgd::|void ", "", name)
    name = re.sub(r"\(.*", "", name)
    return name.replace("(bool)1", "true").replace("(bool)0", "false").replace("(int)", "")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = counts.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P[T\d]+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["total"] += 1
            cur[op] += 1
    names = demangle(list(counts))
    print("SASS mnemonic counts per kernel of libsigsvgd_b200.so (cuobjdump -sass, static counts; sm_100a; final build of round 2).")
    print("UTCHMMA = tcgen05.mma kind::f16, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA load/store,")
    print("UTMAPF = TMA prefetch, SYNCS = mbarrier ops, FFMA2 = packed fp32 FMA, DFMA = fp64 FMA (signature-PDE kernels), MUFU = ex2/rcp/rsqrt,")
    print("F2FP = fp32->fp16x2 pack, REDG = red.global (the K.F chunk reductions, statistics).")
    print()
    print("%-44s" % "kernel" + "%8s" % "total" + "".join("%8s" % c for c in COLS))
    rows = sorted((short(names[k]), v) for k, v in counts.items())
    for name, c in rows:
        print("%-44s" % name[:44] + "%8d" % c["total"] + "".join("%8d" % c[x] for x in COLS))


if __name__ == "__main__":
    sys.exit(main())

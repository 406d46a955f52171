"""SASS opcode histogram per kernel of libqflex_b200.so (cuobjdump -sass): the evidence
that the hot kernels are tcgen05 / TMEM / TMA code (UTCHMMA = tcgen05.mma, LDTM =
tcgen05.ld, UTMALDG = cp.async.bulk.tensor, UTCBAR = tcgen05.commit, SYNCS = mbarrier)
and what the bandwidth-bound ones issue (LDG/STG widths).

    python tools/sass_histogram.py > profiles/sass_histogram_r2.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "qflex_b200", "libqflex_b200.so")
KEY = ("UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "SYNCS",
       "HMMA", "FFMA", "DFMA", "DADD", "LDG", "STG", "LDS", "STS", "SHFL", "BAR", "ATOM", "RED",
       "F2F", "F2FP", "PRMT", "MEMBAR", "CCTL", "ELECT", "USETMAXREG")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), stdout=subprocess.PIPE, text=True).stdout.splitlines()
    print("# cuobjdump -sass qflex_b200/libqflex_b200.so (sm_100a): instructions per kernel, by opcode family")
    print("# full opcode (with modifiers) listed for the tensor-core / TMA / TMEM families")
    for (name, ops), pretty in zip(kernels.items(), demangle):
        total = sum(ops.values())
        fam = collections.Counter()
        detail = collections.Counter()
        for op, n in ops.items():
            base = op.split(".")[0]
            fam[base] += n
            if base in ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UTCCP", "LDG", "STG"):
                detail[op] += n
        pretty = pretty.replace("(anonymous namespace)::", "").replace("void ", "")
        print("\n== %s  (%d instructions)" % (pretty.split("(")[0][:100], total))
        shown = [(k, fam[k]) for k in KEY if fam.get(k)]
        print("   " + "  ".join("%s %d" % kv for kv in shown))
        if detail:
            print("   " + "  ".join("%s x%d" % kv for kv in sorted(detail.items())))


if __name__ == "__main__":
    sys.exit(main())

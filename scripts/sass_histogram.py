"""Per-kernel counts of the SASS instructions that show which hardware path a kernel uses: tcgen05 MMAs (UTCHMMA,
.2CTA = cta_group::2), TMEM loads/stores (LDTM/STTM), TMA tensor loads/stores (UTMALDG/UTMASTG), tcgen05 commits
(UTCBAR), bulk copies (UBLKCP), mbarrier waits (SYNCS), plus the kernel's register count from the ELF.

    python scripts/sass_histogram.py [lib.so] > profiles/r2_sass_histogram.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "alphago_b200", "libalphago_b200.so")
PAT = ["UTCHMMA.2CTA", "UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "UBLKCP", "SYNCS", "HMMA", "IMMA",
       "ATOMS", "ATOMG", "RED", "SHFL", "VOTE", "POPC", "LDG", "STG", "LDS", "STS"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, order, cur = {}, [], None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            order.append(cur)
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        counts[cur]["_total"] += 1
        for p in PAT:
            if op == p or op.startswith(p + "."):
                counts[cur][p] += 1
                break
    names = demangle(order)
    print("SASS instruction histogram of %s (cuobjdump -sass, sm_100a)" % os.path.relpath(LIB, ROOT))
    print("columns: static instruction counts per kernel; UTCHMMA = tcgen05.mma (\".2CTA\" = cta_group::2), LDTM = tcgen05.ld,")
    print("UTMALDG = cp.async.bulk.tensor load, UTCBAR = tcgen05.commit, SYNCS = mbarrier try_wait/arrive\n")
    for f in order:
        c = counts[f]
        short = re.sub(r"\(.*", "", names.get(f, f).replace("(anonymous namespace)::", "").replace("void ", ""))
        keys = [p for p in PAT if c[p]]
        print("%-60s total %6d  %s" % (short[-60:], c["_total"], "  ".join("%s=%d" % (k, c[k]) for k in keys)))


if __name__ == "__main__":
    main()

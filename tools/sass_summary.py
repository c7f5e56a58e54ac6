"""Opcode census of libvisitron_b200.so per kernel (runs without a GPU): what proves which kernels are Blackwell-native.

    python tools/sass_summary.py > profiles/r2_sass_summary.txt

UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG = TMA tensor load, HMMA = mma.sync (legacy tensor path, used
on purpose by the LSTM recurrence only), SYNCS / ... = mbarrier, UCGABAR = cluster barrier, MUFU = special-function unit.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "visitron_b200", "libvisitron_b200.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "HMMA", "LDSM", "SYNCS", "UCGABAR", "LDGSTS", "RED", "ATOM", "MUFU", "STAS"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kern, counts, total = None, collections.OrderedDict(), collections.Counter()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            kern = re.sub(r"\(.*", "", kern)
            counts[kern] = collections.Counter()
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and kern:
            op = m.group(1)
            counts[kern]["_all"] += 1
            for w in WATCH:
                if op.startswith(w):
                    counts[kern][w] += 1
                    total[w] += 1
    print("libvisitron_b200.so: %d kernels; totals: %s" % (len(counts), ", ".join("%s %d" % (w, total[w]) for w in WATCH if total[w])))
    print("%-86s %7s  %s" % ("kernel", "instr", "watched opcodes"))
    for k, c in counts.items():
        w = ", ".join("%s %d" % (x, c[x]) for x in WATCH if c[x])
        print("%-86s %7d  %s" % (k[:86], c["_all"], w))


if __name__ == "__main__":
    main()

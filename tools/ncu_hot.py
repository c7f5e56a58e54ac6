"""Hot instructions of one profiled launch: python tools/ncu_hot.py report.ncu-rep SEGMENT_INDEX|-1 [min_share]  (run here, no GPU)."""
import csv
import io
import subprocess
import sys


def main(rep, idx, min_share=0.006):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    segs = []
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] == "Kernel Name":
            segs.append([])
        if segs:
            segs[-1].append(r)
    if idx < 0:  # list the segments (the source page repeats a kernel once per view)
        for i, sg in enumerate(segs):
            print(i, sg[0][1][:100])
        return
    rd = segs[idx]
    print(rd[0][1][:120])
    hdr, rows = rd[1], [r for r in rd[2:] if len(r) > 10]
    iS, iI, isrc = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[iS]) for r in rows)
    print("total samples", tot, "SASS instructions", len(rows))
    agg = {}
    for r in rows:
        for i in stall_cols:
            agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
    print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / tot) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
    for k, r in enumerate(rows):
        s = int(r[iS])
        if s >= tot * min_share:
            top = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
            print("%5d %6d %5.1f%% x%-8s %-70s %s" % (k, s, 100.0 * s / tot, r[iI], r[isrc].strip()[:70], " ".join("%s:%d" % (n, v) for v, n in top if v)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), float(sys.argv[3]) if len(sys.argv) > 3 else 0.006)

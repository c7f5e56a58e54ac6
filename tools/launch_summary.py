"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (run here, no GPU needed)."""
import collections
import csv
import re
import sys


def main(path, title=""):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    if title:
        print(title)
    print("total %.1f us over %d launches (cold-cache, serialised under ncu: compare shares)" % (tot, sum(c for c, _ in agg.values())))
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-72s %5d %10.1f us %5.1f%%" % (k[:72], c, t, 100 * t / tot))


if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))

"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (run here, no GPU needed).

    python tools/launch_summary.py launches.csv [--step N] [title...]

--step N keeps only the N-th (0-based) step, a step being the launches from one `embed_ln_fwd_kernel`
(the first kernel of the encoder forward) up to the next one.
"""
import collections
import csv
import re
import sys


def main(path, step=None, title=""):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        rows.append((name, v))
    if step is not None:
        starts = [i for i, (n, _) in enumerate(rows) if "embed_ln_fwd_kernel" in n]
        # the step's label bookkeeping (nonzero/index kernels) precedes embed_ln_fwd by a few launches: cut at embed_ln_fwd
        lo = starts[step]
        hi = starts[step + 1] if step + 1 < len(starts) else len(rows)
        rows = rows[lo:hi]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for name, v in rows:
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    if title:
        print(title)
    print("total %.1f us over %d launches (cold-cache, serialised under ncu: compare shares)" % (tot, sum(c for c, _ in agg.values())))
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-72s %5d %10.1f us %5.1f%%" % (k[:72], c, t, 100 * t / tot))


if __name__ == "__main__":
    args = sys.argv[1:]
    step = None
    if "--step" in args:
        i = args.index("--step")
        step = int(args[i + 1])
        del args[i:i + 2]
    main(args[0], step, " ".join(args[1:]))

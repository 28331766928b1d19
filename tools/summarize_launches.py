"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (share of the step).
    python tools/summarize_launches.py profiles/<file>.csv"""
import collections
import csv
import sys


def main(path):
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    agg = collections.defaultdict(list)
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("ns", "nsecond"):
            v /= 1e3
        agg[(r["Kernel Name"].split("(")[0][:48], r["Grid Size"])].append(v)
    tot = sum(sum(v) for v in agg.values())
    print(f"{'kernel':48s} {'grid':>16s} {'n':>5s} {'avg us':>9s} {'total ms':>9s} {'share':>6s}")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k[0]:48s} {k[1]:>16s} {len(v):5d} {sum(v) / len(v):9.1f} {sum(v) / 1e3:9.2f} {sum(v) / tot:6.3f}")
    print(f"total {tot / 1e3:.2f} ms over {sum(len(v) for v in agg.values())} launches")


if __name__ == "__main__":
    main(sys.argv[1])

"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launch count, total device time, share."""
import collections
import csv
import re
import sys


def main(path, top=25):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        ms = v / 1e6 if u.startswith("n") else (v / 1e3 if u.startswith("u") else v)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ms
        n += 1
    tot = sum(a[1] for a in agg.values())
    print(f"{n} launches, {tot:.3f} ms of kernel time (cold-cache, serialised: compare shares)")
    print(f"{'kernel':44s} {'launches':>8s} {'ms':>12s} {'share':>7s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{k:44s} {a[0]:8d} {a[1]:12.3f} {100 * a[1] / tot:6.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)

"""Summarise an ncu CSV holding dram__bytes_read.sum, dram__bytes_write.sum and gpu__time_duration.sum per launch:
per kernel the launch count, mean DRAM bytes per launch and the DRAM rate.  `--json` prints {kernel: mean bytes per launch}."""
import collections
import csv
import json
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6,
        "msecond": 1e-3, "second": 1.0}


def load(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    per = collections.OrderedDict()
    for row in csv.DictReader(lines):
        m = row.get("Metric Name")
        if m not in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"):
            continue
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", "")) * UNIT.get(row["Metric Unit"], 1.0)
        a = per.setdefault(k, {"n": 0, "bytes": 0.0, "s": 0.0})
        if m == "gpu__time_duration.sum":
            a["n"] += 1
            a["s"] += v
        else:
            a["bytes"] += v
    return per


def main():
    path = sys.argv[1]
    per = load(path)
    if "--json" in sys.argv:
        print(json.dumps({k: a["bytes"] / max(a["n"], 1) for k, a in per.items()}, indent=1))
        return
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    tot = sum(a["bytes"] for a in per.values())
    print(f"{sum(a['n'] for a in per.values())} launches, {tot / 1e9:.3f} GB of DRAM traffic")
    print(f"{'kernel':40s} {'launches':>8s} {'MB/launch':>11s} {'total GB':>9s} {'GB/s':>8s}")
    for k, a in sorted(per.items(), key=lambda kv: -kv[1]["bytes"])[:top]:
        print(f"{k:40s} {a['n']:8d} {a['bytes'] / a['n'] / 1e6:11.2f} {a['bytes'] / 1e9:9.3f} {a['bytes'] / max(a['s'], 1e-12) / 1e9:8.0f}")


if __name__ == "__main__":
    main()

"""Per-source-line instruction counts and stall samples from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv --kernel-name regex:K`.
usage: python tools/ncu_lines.py export.csv [ntop]"""
import csv
import sys


def main(path, ntop=25):
    rows = list(csv.reader(open(path)))
    cur_file, hdr, out = None, None, []
    for x in rows:
        if len(x) == 2 and x[0] == "File Path":
            cur_file = x[1].split("/")[-1]
            continue
        if len(x) >= 2 and x[0] == "Line No":
            hdr = {c: i for i, c in enumerate(x)}
            continue
        if hdr is None or len(x) < len(hdr) or x[0] == "":
            continue
        try:
            out.append((cur_file, int(x[0]), x[1].strip(), int(x[hdr["Instructions Executed"]] or 0), int(x[hdr["# Samples"]] or 0),
                        float(x[hdr["Avg. Threads Executed"]] or 0)))
        except ValueError:
            pass
    ti, ts = sum(o[3] for o in out), sum(o[4] for o in out)
    print(f"{path}: {ti / 1e6:.1f} M warp instructions, {ts} stall samples")
    print("-- by instructions executed")
    for o in sorted(out, key=lambda o: -o[3])[:ntop]:
        print(f"{o[0]:>12}:{o[1]:<5} {100 * o[3] / ti:5.1f}% inst {100 * o[4] / max(ts, 1):5.1f}% samp  thr {o[5]:4.1f}  {o[2][:100]}")
    print("-- by stall samples")
    for o in sorted(out, key=lambda o: -o[4])[:ntop]:
        print(f"{o[0]:>12}:{o[1]:<5} {100 * o[3] / ti:5.1f}% inst {100 * o[4] / max(ts, 1):5.1f}% samp  thr {o[5]:4.1f}  {o[2][:100]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)

"""Per-source-line instruction and stall-sample shares from `ncu -i X.ncu-rep --page source --csv
--kernel-name regex:K --print-source cuda,sass > file.csv`."""
import csv
import sys


def main(path, top=40):
    sec, agg = None, {}
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            sec = r[1].split("/")[-1]
            continue
        if not r[0].isdigit():
            continue
        j = next((j for j in range(2, len(r) - 6) if r[j] == "-" and r[j + 1] == "-"), None)
        if j is None:
            continue
        try:
            samples, inst = float(r[j + 4]), float(r[j + 5])
        except ValueError:
            continue
        k = (sec, int(r[0]), ",".join(r[1:j]).strip()[:100])
        a = agg.setdefault(k, [0.0, 0.0])
        a[0] += inst
        a[1] += samples
    tot, tots = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
    print(f"warp instructions {tot:.0f}, stall samples {tots:.0f}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0]}:{k[1]:4d} inst {100 * a[0] / tot:5.1f}% samples {100 * a[1] / max(tots, 1):5.1f}%  {k[2]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)

"""Summarise ncu CSV exports under profiles/: launch list (--metrics gpu__time_duration.sum) -> per-kernel
totals and shares; `--set full` report (--page raw --csv) -> the few metrics the roofline needs."""
import collections
import csv
import re
import sys


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("humid::", "")
        if k.startswith(("at::", "<unnamed>::", "at_cuda", "cub::")):  # torch kernels of the synthetic-data generator
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    out = [f"| kernel | launches | total us | share |", "|---|---|---|---|"]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| `{k}` | {a[0]} | {a[1]:.1f} | {100 * a[1] / tot:.1f} % |")
    out.append(f"| all | {sum(a[0] for a in agg.values())} | {tot:.1f} | 100 % |")
    return "\n".join(out)


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "launch__grid_size", "launch__block_size"]


def full(path):
    r = list(csv.reader(open(path)))
    h = r[0]
    name_col = h.index("Kernel Name")
    out = []
    for row in r[2:]:
        out.append(f"### `{row[name_col].split('(')[0]}`\n")
        out.append("| metric | value | unit |\n|---|---|---|")
        for i, n in enumerate(h):
            if n in WANT:
                out.append(f"| {n} | {row[i]} | {r[1][i]} |")
        out.append("")
    return "\n".join(out)


if __name__ == "__main__":
    print(launches(sys.argv[2]) if sys.argv[1] == "launches" else full(sys.argv[2]))

"""profiles/traffic.json from the CSV of `tools/profile_r02b.sh <tag> traffic`: DRAM bytes (read + write) per stage and
device step = sum over the launches of the stage's kernels / number of finalize calls in the capture (one
table_compact_kernel launch each).  Usage: python tools/traffic_from_ncu.py gpurun_out/<tag>_traffic.csv [--write]"""
import collections
import csv
import json
import re
import sys

STAGE = [  # (regex on the kernel name with its template arguments, stage)
    (r"^table_compact_kernel", "table_compact"), (r"^table_insert_kernel", "table_insert"),
    (r"^radix_pass_kernel<.*SlotDigit", "table_partition"),
    (r"^radix_pass_kernel<.*ShiftDigit, unsigned int>", "sort_words"),
    (r"^radix_pass_kernel<.*ShiftDigit, unsigned long", "join_sort"),
    (r"^radix_digit_hist_kernel<unsigned int>", "sort_words"), (r"^radix_digit_hist_kernel<unsigned long", "join_sort"),
    (r"^gather_sorted_words_kernel|^sort_long_runs_kernel", "rank_columns"),
    (r"^hamming_join_kernel<.*>", "join_hamming"), (r"^csr_fill_kernel", "csr_build"), (r"^uf_union_pairs_kernel", "union_find"),
    (r"^classify_words_kernel", "classify_components"), (r"^replay_small_kernel", "replay_directional"),
]


def main():
    path = sys.argv[1]
    lines = [l for l in open(path) if not l.startswith("==")]
    per_launch = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = re.sub(r"^void ", "", row["Kernel Name"]).replace("humid::", "")
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        if row["Metric Name"].startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            per_launch.setdefault((row["ID"], k), [0.0, 0.0])[0] += v
        elif row["Metric Name"] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1.0)
            per_launch.setdefault((row["ID"], k), [0.0, 0.0])[1] += v
    stage_bytes, stage_us, stage_n = collections.Counter(), collections.Counter(), collections.Counter()
    for (_, k), (b, us) in per_launch.items():
        for rx, st in STAGE:
            if re.search(rx, k):
                stage_bytes[st] += b
                stage_us[st] += us
                stage_n[st] += 1
                break
    steps = max(1, stage_n["table_compact"])
    out = {st: int(stage_bytes[st] / steps) for st in stage_bytes}
    print("finalize calls in the capture:", steps)
    print("| stage | launches per step | DRAM bytes per step | us per step (under ncu: cold, serialised) |\n|---|---|---|---|")
    for st in out:
        print(f"| {st} | {stage_n[st] / steps:.1f} | {out[st]:,} | {stage_us[st] / steps:.0f} |")
    if "--write" in sys.argv:
        p = "profiles/traffic.json"
        d = json.load(open(p))
        d.update(out)
        d["_note_r02q"] = ("table_compact, table_insert, table_partition, sort_words (4 passes over 32-bit prefixes + histogram), "
                           "rank_columns (gather_sorted_words + sort_long_runs), join_sort, csr_build, union_find, classify_components, "
                           "replay_directional re-measured with tools/profile_r02b.sh traffic (dram__bytes_read.sum + dram__bytes_write.sum "
                           "per launch, averaged over the finalize calls of the capture); join_hamming = both block joins")
        json.dump(d, open(p, "w"), indent=1)
        print("updated", p)


if __name__ == "__main__":
    main()

#!/bin/bash
# usage: bash tools/profile_bench.sh <tag>   (on the GPU box, via gpurun): the bench-line part of tools/profile_round.sh —
# the launch list of `bench.py --steps 2 --warmup 3` and one `--set full` capture of the kernels of one device step.
# Each ncu command runs only after the same command line has exited 0 without ncu.
tag=${1:-r01o}
K='regex:^(pack_words|table_|radix_|digit_hist|scan_|gather_columns|scatter_ranks|filter_build|hamming|csr_|uf_|iota_kernel|big_|replay_|compact_clusters|keep_scatter|hist_|run_starts)'
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_s2.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1
echo launches rc=$?
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on \
    -k "regex:^(pack_words|table_insert|radix_pass|gather_columns|hamming_join|csr_fill_kernel|uf_union|replay_kernel)" \
    -s 18 -c 18 -o gpurun_out/${tag}_top python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu2.log 2>&1
echo full rc=$?

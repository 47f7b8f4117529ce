#!/bin/bash
# usage: bash tools/profile_round.sh <tag>   (on the GPU box, via gpurun).  Each ncu command runs only
# after the same command line has exited 0 without ncu.  Outputs go to gpurun_out/<tag>_*.
tag=${1:-r01k}
K='regex:^(pack_words|table_|radix_|digit_hist|scan_|gather_columns|scatter_ranks|filter_build|hamming|csr_|uf_|iota_kernel|big_|replay_|compact_clusters|keep_scatter|hist_|run_starts)'
set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1
echo launches rc=$?
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 || exit 1
# second device step of that command: 1 pack, 2 table_insert-related, 10 radix passes, gather, 2 joins, csr_fill, uf_union, replay
ncu --set full --clock-control none --import-source on \
    -k "regex:^(pack_words|table_insert|radix_pass|gather_columns|hamming_join|csr_fill_kernel|uf_union|replay_kernel)" \
    -s 18 -c 18 -o gpurun_out/${tag}_top python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu2.log 2>&1
echo full rc=$?
# the other paths: edit distance 2 (configs[3]), Hamming 2 by block pairs (configs[2]), words of 36 letters
python tools/stage_table.py 3 20000000 1 > gpurun_out/${tag}_stages_c4.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k "regex:^(edit_replay_tasks|deletion_join|deletion_variants|edit_replay_single)" -c 4 \
    -o gpurun_out/${tag}_edit python tools/stage_table.py 3 20000000 1 > gpurun_out/${tag}_ncu3.log 2>&1
echo edit rc=$?
python tools/stage_table.py 1 50000000 1 36 > gpurun_out/${tag}_stages_n36.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k "regex:^(hamming_join_wide|wide_|gather_u64|gather_halves)" -c 8 \
    -o gpurun_out/${tag}_wide python tools/stage_table.py 1 50000000 1 36 > gpurun_out/${tag}_ncu4.log 2>&1
echo wide rc=$?
python tools/stage_table.py 2 50000000 1 > gpurun_out/${tag}_stages_c3.log 2>&1

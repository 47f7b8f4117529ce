#!/bin/bash
# usage (via gpurun, one GPU): bash tools/profile_r02.sh <tag>.  Each ncu command runs only after the same command line
# has exited 0 without ncu.  Outputs: gpurun_out/<tag>_*.
tag=${1:-r02p}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cli"
$CMD > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_bench.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
CMD2="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-cli"
$CMD2 > /dev/null 2>&1 || { echo "plain run 2 failed"; exit 1; }
# second device step: pack 1, table_insert 1, radix_pass 10, gather_columns 1, hamming_join 2, csr_fill 1, uf_union_pairs 1,
# classify_words 1, replay_small 1 = 19 launches of these kernels per step
ncu --set full --clock-control none --import-source on \
    -k "regex:^(pack_words_kernel|table_insert|radix_pass|gather_columns|hamming_join|csr_fill_kernel|uf_union_pairs|classify_words|replay_small)" \
    -s 19 -c 19 -o gpurun_out/${tag}_top $CMD2 > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/${tag}_*

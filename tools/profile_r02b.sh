#!/bin/bash
# usage (via gpurun, one GPU, ONE ncu run per call): bash tools/profile_r02b.sh <tag> list|traffic
#   list     the launch list of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cli` (gpu__time_duration.sum)
#   traffic  DRAM bytes + duration of every launch of the stage kernels in `python bench.py --steps 1 --warmup 1 ...`
# The ncu command runs only after the same command line has exited 0 without ncu.  Outputs: gpurun_out/<tag>_*.
tag=${1:-r02q}
mode=${2:-list}
mkdir -p gpurun_out
if [ "$mode" = list ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cli"
  $CMD > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err || { echo "plain run failed"; tail -5 gpurun_out/${tag}_bench.err; exit 1; }
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
  echo "launch list rc=$?"
else
  CMD2="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-cli"
  $CMD2 > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
      -k "regex:^(table_compact|table_insert|radix_pass|radix_digit_hist|gather_sorted_words|sort_long_runs|hamming_join|csr_fill_kernel|uf_union_pairs|classify_words|replay_small)" \
      --log-file gpurun_out/${tag}_traffic.csv $CMD2 > gpurun_out/${tag}_ncu2.log 2>&1
  echo "traffic capture rc=$?"
fi
ls -la gpurun_out/${tag}_*

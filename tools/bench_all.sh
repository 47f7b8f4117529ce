#!/bin/bash
# usage (via gpurun): bash tools/bench_all.sh <tag> — one bench line per BASELINE.json configuration that fits one GPU
tag=${1:-all}
mkdir -p gpurun_out
for spec in "0 1000000" "1 50000000" "2 50000000" "3 20000000"; do
  set -- $spec
  timeout 900 python bench.py --config $1 --tuples $2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_c$1.log 2>&1
  echo "config $1 rc=$?"; python tools/bench_brief.py < gpurun_out/${tag}_bench_c$1.log
done
timeout 900 python bench.py --config 1 --tuples 50000000 --word-length 36 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_c1_n36.log 2>&1
echo "config 1 -n 36 rc=$?"; python tools/bench_brief.py < gpurun_out/${tag}_bench_c1_n36.log
timeout 1200 python bench.py --config 4 --tuples 500000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_c4.log 2>&1
echo "config 4 rc=$?"; python tools/bench_brief.py < gpurun_out/${tag}_bench_c4.log

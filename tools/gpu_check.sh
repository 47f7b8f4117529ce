#!/bin/bash
# usage (via gpurun): bash tools/gpu_check.sh <tag> — new tests first, then the whole GPU suite, stage tables, bench
tag=${1:-check}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -k "wide or n36 or n50" -x > gpurun_out/${tag}_wide.log 2>&1; echo "wide rc=$?"
tail -5 gpurun_out/${tag}_wide.log
timeout 1500 python -m pytest tests -q -m gpu --durations=15 > gpurun_out/${tag}_gpu_tests.log 2>&1; echo "all rc=$?"
tail -25 gpurun_out/${tag}_gpu_tests.log
timeout 300 python tools/stage_table.py 1 50000000 3 36 > gpurun_out/${tag}_stages_n36.log 2>&1; echo "n36 rc=$?"
cat gpurun_out/${tag}_stages_n36.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.log 2>&1; echo "bench rc=$?"
python tools/bench_brief.py < gpurun_out/${tag}_bench.log
timeout 600 python bench.py --steps 3 --warmup 3 --word-length 36 --no-cpu-baseline > gpurun_out/${tag}_bench_n36.log 2>&1; echo "bench36 rc=$?"
python tools/bench_brief.py < gpurun_out/${tag}_bench_n36.log

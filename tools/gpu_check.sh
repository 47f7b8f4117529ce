#!/bin/bash
# usage (via gpurun): bash tools/gpu_check.sh <tag> — the whole GPU suite, smoke, the bench line, the -n 36 bench
tag=${1:-check}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --durations=8 > gpurun_out/${tag}_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -14 gpurun_out/${tag}_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.log 2>&1; echo "bench rc=$?"
python tools/bench_brief.py < gpurun_out/${tag}_bench.log
timeout 600 python bench.py --steps 3 --warmup 3 --word-length 36 --no-cpu-baseline > gpurun_out/${tag}_bench_n36.log 2>&1; echo "bench36 rc=$?"
python tools/bench_brief.py < gpurun_out/${tag}_bench_n36.log

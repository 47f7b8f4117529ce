#!/bin/bash
# usage (via gpurun): bash tools/r02_check.sh <tag> [cli-scale] — GPU suite, smoke, both bench arms (reference first,
# so that the FastQ files of the whole-program legs are shared)
tag=${1:-r02a}
scale=${2:-0.25}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --durations=8 > gpurun_out/${tag}_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -14 gpurun_out/${tag}_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 --cli-scale $scale > gpurun_out/${tag}_bench_ref.log 2>gpurun_out/${tag}_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 5 --warmup 3 --cli-scale $scale > gpurun_out/${tag}_bench.log 2>gpurun_out/${tag}_bench.err; echo "bench rc=$?"
tail -5 gpurun_out/${tag}_bench.err
python tools/bench_brief.py < gpurun_out/${tag}_bench.log

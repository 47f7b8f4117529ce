#!/bin/bash
# usage (via gpurun --gpus 8): bash tools/scale_bench.sh <tag> "<N list>" [steps] — bench.py at each N, phases and brief lines
tag=${1:-scale}; ns=${2:-"8 4 2"}; steps=${3:-4}
mkdir -p gpurun_out
for n in $ns; do
  timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps $steps --warmup 3 --no-cli > gpurun_out/${tag}_bench$n.log 2>gpurun_out/${tag}_bench$n.err
  echo "bench N=$n rc=$?"
  grep -a "phases\|Error\|error\|falling" gpurun_out/${tag}_bench$n.err | tail -4
  python tools/bench_brief.py < gpurun_out/${tag}_bench$n.log 2>/dev/null | sed -n '1p;3p'
  python -c "
import json
for l in open('gpurun_out/${tag}_bench$n.log'):
    if l.startswith('{'):
        d = json.loads(l); print('parity equal:', (d.get('parity') or {}).get('equal'))
"
done

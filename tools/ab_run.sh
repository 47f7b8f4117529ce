#!/bin/bash
# usage (on the GPU box, via gpurun): bash tools/ab_run.sh <config> <tuples> <word length or ""> <name> [<name> ...]
# Per-stage device times (tools/stage_table.py) of the default library and of every ab/lib_<name>.so built by
# tools/ab_build.py: one line per library, the result counts first (they must agree), then "stage ms | ...".
ci=$1; R=$2; n=$3; shift 3
for v in "" "$@"; do
  printf "%-14s" "${v:-default}"
  HUMID_GPU_LIB=${v:+$PWD/ab/lib_$v.so} timeout 120 python tools/stage_table.py $ci $R 4 $n 2>&1 |
    grep -E " ms |rror|clusters'" | awk '{if ($0 ~ /clusters/) printf "%s ", $NF; else if ($0 ~ / ms /) printf "%s %s | ", $1, $2; else print} END {print ""}'
done

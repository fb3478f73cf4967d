#!/bin/bash
# A/B of two builds of the library on one box: tools/ab.sh <tag> <variant.so> [bench args...]
TAG=$1; VAR=$2; shift 2
mkdir -p gpurun_out/$TAG
timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" > gpurun_out/$TAG/bench_A.json 2> gpurun_out/$TAG/bench_A.err; echo "A rc=$?"
cp memsurfer_b200/libmemsurfer_b200.so /tmp/libA.so; cp "$VAR" memsurfer_b200/libmemsurfer_b200.so
timeout 300 python bench.py --no-cpu-baseline --no-e2e "$@" > gpurun_out/$TAG/bench_B.json 2> gpurun_out/$TAG/bench_B.err; echo "B rc=$?"
cp /tmp/libA.so memsurfer_b200/libmemsurfer_b200.so

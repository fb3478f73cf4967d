#!/bin/bash
# usage (on the GPU box, from the repo root): tools/gpu_suite.sh <tag> [bench args...]   -> gpurun_out/<tag>/{suite.log,bench.json}
tag=$1; shift
mkdir -p gpurun_out/$tag
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 1200 2>&1 | tail -60 > gpurun_out/$tag/suite.log
tail -15 gpurun_out/$tag/suite.log
if [ "$1" != "nobench" ]; then
  timeout 900 python bench.py "$@" > gpurun_out/$tag/bench.json 2> gpurun_out/$tag/bench.err
  tail -3 gpurun_out/$tag/bench.err
  python tools/show_bench.py gpurun_out/$tag/bench.json 2>/dev/null | head -40
fi

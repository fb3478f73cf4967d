#!/bin/bash
# Multi-GPU visit: tools/gpu_scale.sh <tag> <N>  -- the bench line at N ranks (weak scaling) and the reference arm under torchrun.
TAG=${1:-scale}; N=${2:-2}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/smi.txt 2>&1
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err; echo "bench$N rc=$?" | tee -a $O/rc.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $O/bench_ref_${N}gpu.json 2> $O/bench_ref_${N}gpu.err; echo "ref$N rc=$?" | tee -a $O/rc.txt
cut -c1-300 $O/bench_${N}gpu.json

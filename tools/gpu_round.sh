#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, the ncu launch list of the bench command and
# `--set full` captures of the kernels that dominate a frame.  Everything lands in gpurun_out/<tag>/.
# usage: tools/gpu_round.sh <tag> [tests|bench|launches|ncu ...]   (default: all four)
set -u
TAG=${1:-run}; shift || true
WHAT=${*:-tests bench launches ncu}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/smi.txt" 2>&1
for w in $WHAT; do
case $w in
tests)
  timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$OUT/rc.txt"
  timeout 300 python __graft_entry__.py smoke > "$OUT/smoke.log" 2>&1; echo "smoke rc=$?" | tee -a "$OUT/rc.txt" ;;
bench)
  timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > "$OUT/bench_ref.json" 2> "$OUT/bench_ref.err"; echo "bench_ref rc=$?" | tee -a "$OUT/rc.txt"
  timeout 600 python bench.py > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench rc=$?" | tee -a "$OUT/rc.txt"
  ;;
launches)
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file "$OUT/launches.csv" \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-extras --single-stream > "$OUT/ncu_launches.log" 2>&1; echo "launches rc=$?" | tee -a "$OUT/rc.txt" ;;
ncu)
  for spec in "labels:k_kde_groups:4" "mesh:k_pad_verts|k_corner_terms|k_vertex_gather8:3" "normals:k_knn_select|k_knn_thread|k_pca|k_orient:3"; do
    mode=${spec%%:*}; rest=${spec#*:}; rx=${rest%%:*}; cnt=${rest##*:}
    timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$rx" -c "$cnt" -f -o "$OUT/ncu_$mode" \
        python tools/prof_kde.py --mode "$mode" > "$OUT/ncu_$mode.log" 2>&1; echo "ncu_$mode rc=$?" | tee -a "$OUT/rc.txt"
  done ;;
esac
done
cat "$OUT/rc.txt"
tail -3 "$OUT/pytest_gpu.log" 2>/dev/null
cat "$OUT/bench.json" 2>/dev/null | cut -c1-600

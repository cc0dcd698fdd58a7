#!/bin/bash
# one round-check call: GPU tests, both bench arms, ncu launch list, ncu --set full of the hot kernels
# usage: tools/round_gpu.sh <tag> [skip-tests]
# afterwards, here: python tools/ncu_traffic.py gpurun_out/prof_<tag>.ncu-rep 16777216   (rewrites
# profiles/ncu_traffic.json, tagged with the kernel-source hash bench.py checks)
tag=${1:-vX}
o=gpurun_out
mkdir -p $o
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > $o/gpu_$tag.txt
if [ "$2" != "skip-tests" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > $o/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $o/pytest_gpu_$tag.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 > $o/bench_$tag.json 2> $o/bench_$tag.err; echo "bench rc=$?"; cut -c1-400 $o/bench_$tag.json
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > $o/bench_ref_$tag.json 2> $o/bench_ref_$tag.err; echo "ref rc=$?"; cut -c1-300 $o/bench_ref_$tag.json
short="python bench.py --steps 2 --warmup 3 --no-cpu --no-secondary --e2e-steps 0"
$short > $o/plain_$tag.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches_$tag.csv \
  $short > $o/ncu_launch_$tag.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_density_union|k_forces_union|k_advect|k_reorder|k_reference_rank|k_morton_hist|k_place|k_locate_count|k_order_cells' \
  --launch-skip 24 --launch-count 10 -o $o/prof_$tag -f \
  $short > $o/ncu_full_$tag.log 2>&1; echo "ncu full rc=$?"
ls -la $o | tail -20

#!/bin/bash
# Full GPU-box pass: tests, smoke, bench (both arms), ncu launch list, ncu --set full of the heavy kernels.
# gpurun --timeout 1500 -- bash scripts/gpu_full.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
bash scripts/gpu_round.sh $tag 30
timeout 120 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo smoke_exit=$?; tail -1 gpurun_out/${tag}_smoke.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.log 2>&1; echo ref_exit=$?; tail -1 gpurun_out/${tag}_bench_ref.log | cut -c1-400
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_launches.log 2>&1; echo ncu_list_exit=$?
timeout 500 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|bucket|bin_|sort|emit|gather|scan' \
  --launch-skip 60 --launch-count 16 -o gpurun_out/${tag}_full -f \
  python bench.py --steps 3 --warmup 8 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu_full.log 2>&1; echo ncu_full_exit=$?
ls -la gpurun_out | tail -12

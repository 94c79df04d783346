#!/bin/bash
# Standard GPU-box check: `gpurun -- bash scripts/gpu_round.sh <tag> [steps]` -> gpurun_out/<tag>_*.log
tag=${1:-rX}; steps=${2:-30}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo pytest_exit=$?
tail -4 gpurun_out/${tag}_pytest_gpu.log
timeout 400 python bench.py --steps $steps --warmup 5 > gpurun_out/${tag}_bench.log 2>&1; echo bench_exit=$?
tail -1 gpurun_out/${tag}_bench.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('value', d['value'], 'ms/step', d['ms_per_step'], 'e2e', d['e2e'], 'host', d.get('host_wall_ms_per_step'))
    print('stages', d['roofline']['stage_ms'])
    print('roofline', {k: d['roofline'][k] for k in ('kernel', 'achieved', 'frac')}, 'cpu', d.get('cpu_baseline', {}).get('value'))
except Exception as e:
    print('no bench line', e)
"

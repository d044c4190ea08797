#!/bin/bash
# One GPU-box pass: parity tests in separate processes (a faulting kernel must not poison the rest), smoke, bench.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_check.sh [quick]
set -u
mkdir -p gpurun_out
PY="python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout=300 -p no:cacheprovider"
run() { name=$1; shift; echo "=== $name"; timeout 1200 "$@" > gpurun_out/$name.log 2>&1; echo "exit $?" | tee -a gpurun_out/$name.log; tail -n 25 gpurun_out/$name.log; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv | tee gpurun_out/gpu.txt
run t_simt    $PY -k "fp32 or lookup_on_golden or rejects or alt_corr"
run t_homog   $PY -k "homography"
run t_tc      $PY -k "tf32 or bf16"
run t_net     $PY -k "wrapper or tracker"
run smoke     python -c "import __graft_entry__ as g; g.smoke()"
if [ "${1:-}" != "quick" ]; then
  run bench   python bench.py --steps 10 --warmup 3
fi

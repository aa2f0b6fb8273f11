#!/bin/bash
# kernel parity after a k_sweep change (replay of every decision, cull-disabled bit identity, walls), then the bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "sweep_replay or cull_never or test_gpu_walls or test_gpu_parity or test_gpu_errors or test_gpu_observables" > gpurun_out/pytest_kernel.log 2>&1; echo "pytest(kernel) exit $?"
tail -4 gpurun_out/pytest_kernel.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
b=json.load(open('gpurun_out/bench.json'))
print({k:b[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', b['e2e']['value'], 'roofline', b['roofline']['frac'], b['roofline']['avg_launch_ms'], b['breakdown_ms_per_step'], b['clocks'])
PY
tail -3 gpurun_out/bench.err

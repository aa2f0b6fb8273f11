#!/bin/bash
# one-off: kernel parity after a k_sweep change, then the bench, then one statistics case
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "sweep_replay or cull_never or test_gpu_walls" > gpurun_out/pytest_kernel.log 2>&1; echo "pytest(kernel) exit $?"
tail -4 gpurun_out/pytest_kernel.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 1500 python -m pytest tests/test_gpu_statistics.py -m gpu -q -s -k "${1:-sc_l5_isotropic}" > gpurun_out/pytest_stat.log 2>&1; echo "pytest(stat) exit $?"
grep -n "^\[\|  density\|  g_shell\|  p2\|  tau\|passed\|failed\|^E  " gpurun_out/pytest_stat.log | tail -30

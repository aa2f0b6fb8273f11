#!/bin/bash
# GPU parity run: full -m gpu suite (statistics printed), then the contract bench; PAIRS=1e8 adds the 10^8-pair run
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -k "statistics or errors" > gpurun_out/pytest_stat.log 2>&1; echo "pytest(stat) exit $?" >> gpurun_out/pytest_stat.log
grep -n "^\[\|  density\|  g_shell\|  p2\|  tau\|passed\|failed\|exit" gpurun_out/pytest_stat.log | tail -40
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_statistics.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
if [ -n "$PAIRS" ]; then
timeout 1500 python tools/pairs_1e8.py --target $PAIRS > gpurun_out/pairs_1e8.json 2> gpurun_out/pairs_1e8.err; echo "pairs exit $?"
tail -3 gpurun_out/pairs_1e8.err; head -c 1500 gpurun_out/pairs_1e8.json; echo
fi
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err

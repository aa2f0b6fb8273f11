#!/bin/bash
# small-grid rounding change: whole GPU suite, then the default bench line (all configs)
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -x -m gpu -q > gpurun_out/sg_pytest.log 2>&1 ) 2>&1 | grep real >> gpurun_out/sg_pytest.log; tail -4 gpurun_out/sg_pytest.log
( time timeout 1500 python bench.py > gpurun_out/sg_bench.json 2> gpurun_out/sg_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/sg_bench.err
timeout 300 python __graft_entry__.py --smoke > gpurun_out/sg_smoke.log 2>&1; tail -1 gpurun_out/sg_smoke.log

#!/bin/bash
# whole GPU suite, then the default bench line (all configs), then smoke
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -x -m gpu -q > gpurun_out/sb_pytest.log 2>&1 ) 2>&1 | grep real >> gpurun_out/sb_pytest.log; tail -4 gpurun_out/sb_pytest.log
( time timeout 1500 python bench.py > gpurun_out/sb_bench.json 2> gpurun_out/sb_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/sb_bench.err
timeout 300 python __graft_entry__.py --smoke > gpurun_out/sb_smoke.log 2>&1; tail -1 gpurun_out/sb_smoke.log

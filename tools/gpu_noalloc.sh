#!/bin/bash
# the no-allocation entry point's test alone in a fresh process, then the whole GPU suite (it collects last), then smoke
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
( time timeout 2400 python -m pytest tests -x -m gpu -q > gpurun_out/noalloc_pytest.log 2>&1 ) 2>&1 | grep real >> gpurun_out/noalloc_pytest.log; tail -4 gpurun_out/noalloc_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/noalloc_smoke.log 2>&1; tail -1 gpurun_out/noalloc_smoke.log

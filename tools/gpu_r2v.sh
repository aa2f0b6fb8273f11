#!/bin/bash
# round 2, call V: in-cell pair items interleaved across warps (generic sweep) — parity + timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_xc.py tests/test_gpu_sweep_replay.py tests/test_gpu_walls.py tests/test_gpu_relaxation.py -q 2>&1 | tail -5
timeout 600 python tools/quick_gen.py convex_200k kmer4_200k kmer3_wca_500k 2>&1 | tail -3
COUNTING=1 timeout 600 python tools/quick_gen.py convex_200k kmer4_200k 2>&1 | tail -2

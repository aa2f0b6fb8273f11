#!/bin/bash
# round 2, call W: register cap for the XenoCollide instantiation of the generic sweep
timeout 600 python -m pytest tests/test_gpu_xc.py -q 2>&1 | tail -2
timeout 600 python tools/quick_gen.py convex_200k 2>&1 | tail -1
COUNTING=1 timeout 600 python tools/quick_gen.py convex_200k 2>&1 | tail -1

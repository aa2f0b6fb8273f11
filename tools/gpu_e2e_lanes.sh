#!/bin/bash
# cycles with an asynchronous snapshot per cycle (s) against cycles with a synchronous download per cycle (d)
mkdir -p gpurun_out
for a in "1 cs" "1 cd" "2 cs"; do set -- $a; timeout 60 python tools/e2e_probe.py $1 $2 2>&1 | tail -$(( $1 + 1 )); done > gpurun_out/e2e_snapshot.txt 2>&1
cat gpurun_out/e2e_snapshot.txt

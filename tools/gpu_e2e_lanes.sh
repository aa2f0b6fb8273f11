#!/bin/bash
# e2e pipelining probe: packings in flight per GPU, phases separately
mkdir -p gpurun_out
for a in "1 c" "1 ucd" "3 ucd" "4 ucd" "4 cd" "4 uc" "4 c" "6 ucd"; do set -- $a; timeout 200 python tools/e2e_probe.py $1 $2 2>&1 | tail -$(( $1 + 1 )); done > gpurun_out/e2e_modes2.txt 2>&1
cat gpurun_out/e2e_modes2.txt

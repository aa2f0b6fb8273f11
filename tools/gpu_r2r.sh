#!/bin/bash
# round 2, call R: parity suite, the ordered-state statistics, generic timings after the fused soft pair function / small cells
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -k "not statistics" 2>&1 | tail -6
( time timeout 1500 python -m pytest tests/test_gpu_statistics.py -q -s -k "sc_l5_nematic or sc_l5_smectic" > gpurun_out/r2r_stat.log 2>&1 ) 2>&1 | grep real
grep -v "^$" gpurun_out/r2r_stat.log | tail -25
timeout 900 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k 2>&1 | tail -4
COUNTING=1 timeout 600 python tools/quick_gen.py kmer4_200k 2>&1 | tail -1
timeout 1200 python bench.py --configs kmer3_wca_500k,kmer3_lj_500k,kmer4_200k,convex_200k --no-cpu-baseline --therm 100 --steps 5 --warmup 3 > gpurun_out/r2r_bench_gen.json 2> gpurun_out/r2r_bench_gen.err; tail -3 gpurun_out/r2r_bench_gen.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2r_bench_gen.json'))
for k,v in d['configs'].items():
    print(k, {x:v.get(x) for x in ('value','ms_per_step','box_moves_per_s','overlap_relaxation','counting_mode','acceptance')}, v['roofline'].get('sweep_ms_per_step'))
PY

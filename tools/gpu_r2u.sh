#!/bin/bash
# round 2, call U: the whole statistics suite with concurrent seeds (duration!), hs_1k_gas with 512-thread CTAs
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests/test_gpu_statistics.py -q -s > gpurun_out/r2u_stat.log 2>&1 ) 2>&1 | grep real
grep -v "^$" gpurun_out/r2u_stat.log | tail -60
timeout 600 python bench.py --configs hs_1k_gas,hs_1m_liquid --no-cpu-baseline --therm 50 --steps 5 --warmup 3 > gpurun_out/r2u_bench.json 2> gpurun_out/r2u_bench.err; tail -3 gpurun_out/r2u_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2u_bench.json'))
for k,v in d['configs'].items():
    print(k, {x:v.get(x) for x in ('value','ms_per_step','box_moves_per_s','acceptance')}, v['roofline'].get('sweep_ms_per_step'))
PY

#!/bin/bash
# round 2, call L: parity after the LJ / recount changes, generic timings, ordered snapshots for the statistics test
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "not statistics" 2>&1 | tail -5
timeout 900 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k > gpurun_out/r2l_quick.log 2>&1; cat gpurun_out/r2l_quick.log
timeout 900 python bench.py --configs kmer3_wca_500k,kmer3_lj_500k,kmer4_200k --no-cpu-baseline --therm 100 --steps 5 --warmup 3 > gpurun_out/r2l_bench_gen.json 2> gpurun_out/r2l_bench_gen.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2l_bench_gen.json'))
for k,v in d['configs'].items():
    print(k, {x:v.get(x) for x in ('value','ms_per_step','box_moves_per_s','overlap_relaxation')}, v['roofline'].get('sweep_ms_per_step'))
PY
EQ_CYCLES=400000 timeout 1500 python tests/golden/make_ordered_snapshots.py > gpurun_out/r2l_ordered.log 2>&1; tail -45 gpurun_out/r2l_ordered.log

#!/bin/bash
# round 2, call F: the full default bench line (with the other BASELINE configs) and the reference arm
mkdir -p gpurun_out
( time timeout 1200 python bench.py > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err ) 2>&1 | grep real; echo "bench exit $?"; tail -5 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2f_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','acceptance','step_sizes','volume_moves','breakdown_ms_per_step','clocks')})
print(d['e2e']); print(d['roofline']['frac'], d['roofline']['physical'])
for k,v in (d.get('configs') or {}).items():
    print(k, json.dumps(v)[:900])
PY
( time timeout 1200 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2f_bench_ref.json 2> gpurun_out/r2f_bench_ref.err ) 2>&1 | grep real; cat gpurun_out/r2f_bench_ref.json | cut -c1-1200; tail -3 gpurun_out/r2f_bench_ref.err

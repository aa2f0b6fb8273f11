#!/bin/bash
# round-end evidence run (1 GPU): full GPU test suite, smoke, the default bench line + reference arm, ncu launch list of
# the bench command
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -x -m gpu -q > gpurun_out/final_pytest.log 2>&1 ) 2>&1 | grep real >> gpurun_out/final_pytest.log; tail -4 gpurun_out/final_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/final_smoke.log 2>&1; tail -1 gpurun_out/final_smoke.log
( time timeout 1500 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/final_bench.err
( time timeout 1200 python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err ) 2>&1 | grep real; cut -c1-300 gpurun_out/final_bench_ref.json
B="python bench.py --configs none --no-cpu-baseline --therm 300 --steps 3 --warmup 3"
E2E_PIPELINE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/final_launches_all.csv $B > gpurun_out/final_ncu_list.log 2>&1
echo "ncu list exit $?"
python - <<'PY'
# keep the 3 warm-up + 3 timed cycles: bench.py downloads the state once after the thermalisation and once after the
# timed steps, so they are the launches between the first two k_pack_shapes
raw = open('gpurun_out/final_launches_all.csv').read().split('\n')
h = [i for i, r in enumerate(raw) if r.startswith('"ID"')][0]
rows = raw[h + 1:]
pack = [i for i, r in enumerate(rows) if 'k_pack_shapes' in r]
open('gpurun_out/final_launches_bench_sc1m.csv', 'w').write('\n'.join([raw[h]] + rows[pack[0] + 1:pack[1]]) + '\n')
PY
rm -f gpurun_out/final_launches_all.csv
ls -la gpurun_out | head -30

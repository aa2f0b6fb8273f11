#!/bin/bash
# round-end evidence run (1 GPU): full GPU test suite, smoke, the default bench line + reference arm, ncu launch list of
# the bench command, ncu --set full summaries of k_gen_sweep per shape (reports are summarised here: gpurun_out/ is
# limited to 64 MiB)
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/final_pytest.log 2>&1 ) 2>&1 | grep real >> gpurun_out/final_pytest.log; tail -4 gpurun_out/final_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/final_smoke.log 2>&1; tail -1 gpurun_out/final_smoke.log
( time timeout 1500 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/final_bench.err
( time timeout 1200 python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err ) 2>&1 | grep real; cut -c1-600 gpurun_out/final_bench_ref.json
B="python bench.py --configs none --no-cpu-baseline --therm 300 --steps 3 --warmup 3"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/final_launches_all.csv $B > gpurun_out/final_ncu_list.log 2>&1
echo "ncu list exit $?"; grep -n "k_sweep_report" gpurun_out/final_launches_all.csv | tail -8 | head -1
python - <<'PY'
# keep the launches of the warm-up + timed steps (the last 6 cycles before the end-to-end section) — ~330 rows
rows = open('gpurun_out/final_launches_all.csv').read().split('\n')
hdr = [i for i, r in enumerate(rows) if r.startswith('"ID"')]
idx = [i for i, r in enumerate(rows) if 'k_sweep_report' in r]
start = idx[-12] if len(idx) >= 12 else 0
open('gpurun_out/final_launches_bench_sc1m.csv', 'w').write('\n'.join([rows[hdr[0]]] + rows[start:start + 420] if hdr else rows[-420:]))
PY
rm -f gpurun_out/final_launches_all.csv
for w in kmer3_lj_500k kmer3_wca_500k kmer4_200k convex_200k; do
  THERM=2 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_gen_sweep" -s 17 -c 1 -o gpurun_out/final_prof_gen_$w -f python tools/quick_gen.py $w > gpurun_out/final_ncu_gen_$w.log 2>&1
  python tools/ncu_summary.py gpurun_out/final_prof_gen_$w.ncu-rep 25 > gpurun_out/final_ncu_gen_$w.txt 2>&1
  python tools/ncu_lines.py gpurun_out/final_prof_gen_$w.ncu-rep 0 40 >> gpurun_out/final_ncu_gen_$w.txt 2>&1
  rm -f gpurun_out/final_prof_gen_$w.ncu-rep
done
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_recount_cull$" -s 12 -c 1 -o gpurun_out/final_prof_recount -f python tools/quick_bench.py 1M > gpurun_out/final_ncu_recount.log 2>&1
python tools/ncu_summary.py gpurun_out/final_prof_recount.ncu-rep 25 > gpurun_out/final_ncu_recount.txt 2>&1
rm -f gpurun_out/final_prof_recount.ncu-rep
ls -la gpurun_out | head -40

#!/bin/bash
# round 2, call K: launch list of the thermalised sc_1m step, ncu --set full of k_sweep on that state and of k_gen_sweep per shape
# (reports are summarised on the box: gpurun_out/ is limited to 64 MiB)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "not statistics" 2>&1 | tail -3
B="python bench.py --configs none --no-cpu-baseline --therm 300 --steps 3 --warmup 3"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2k_launches_all.csv $B > gpurun_out/r2k_ncu_list.log 2>&1
echo "ncu list exit $?"; tail -n 400 gpurun_out/r2k_launches_all.csv > gpurun_out/r2k_launches_bench_sc1m.csv; rm -f gpurun_out/r2k_launches_all.csv
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep<" -s 2410 -c 1 -o gpurun_out/r2k_prof_k_sweep -f $B > gpurun_out/r2k_ncu_full.log 2>&1
echo "ncu full exit $?"
python tools/ncu_summary.py gpurun_out/r2k_prof_k_sweep.ncu-rep 40 > gpurun_out/r2k_ncu_k_sweep.txt 2>&1
python tools/ncu_lines.py gpurun_out/r2k_prof_k_sweep.ncu-rep >> gpurun_out/r2k_ncu_k_sweep.txt 2>&1
for w in kmer3_lj_500k kmer3_wca_500k kmer4_200k convex_200k; do
  THERM=2 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_gen_sweep" -s 17 -c 1 -o gpurun_out/r2k_prof_gen_$w -f python tools/quick_gen.py $w > gpurun_out/r2k_ncu_gen_$w.log 2>&1
  echo "ncu gen $w exit $?"
  python tools/ncu_summary.py gpurun_out/r2k_prof_gen_$w.ncu-rep 40 > gpurun_out/r2k_ncu_gen_$w.txt 2>&1
  python tools/ncu_lines.py gpurun_out/r2k_prof_gen_$w.ncu-rep >> gpurun_out/r2k_ncu_gen_$w.txt 2>&1
  rm -f gpurun_out/r2k_prof_gen_$w.ncu-rep
done
for w in kmer3_lj_500k kmer3_wca_500k kmer4_200k; do
THERM=2 REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2k_launches_gen_$w.csv python tools/quick_gen.py $w > /dev/null 2>&1
done
timeout 900 python bench.py --configs kmer3_wca_500k,kmer4_200k --no-cpu-baseline --therm 100 --steps 5 --warmup 3 > gpurun_out/r2k_bench_gen.json 2> gpurun_out/r2k_bench_gen.err
ls -la gpurun_out

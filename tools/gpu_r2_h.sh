#!/bin/bash
# round-2 GPU call H (1 GPU): bench lines for profiles/, ncu launch list + full captures (raw CSV pages), compute-sanitizer
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== bench c2"; timeout 600 python bench.py --steps 200 --warmup 5 > $O/r2_bench_c2.json 2> $O/r2_bench_c2.err; echo rc=$?
echo "== bench c4 1 gpu"; timeout 600 python bench.py --workload c4 --steps 5 --warmup 3 --no-cpu --no-extras > $O/r2_bench_c4_1gpu.json 2> $O/r2_bench_c4.err; echo rc=$?; cut -c1-300 $O/r2_bench_c4_1gpu.json
echo "== bench c5 quant"; timeout 600 python bench.py --workload c5 --steps 20 --warmup 3 --no-cpu --no-extras > $O/r2_bench_c5_quant_1gpu.json 2> $O/r2_bench_c5.err; echo rc=$?
echo "== launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-extras > $O/ncu_launch.log 2>&1; echo rc=$?
cap() { # name, skip, count, kernel regex, env..., cmd
  name=$1; skip=$2; cnt=$3; pat=$4; shift 4
  timeout 900 env "$@" ncu --set full --clock-control none -k regex:$pat -s $skip -c $cnt -o $O/$name > $O/$name.log 2>&1; echo "$name rc=$?"
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null; rm -f $O/$name.ncu-rep
}
echo "== ncu full"
cap r2_ncu_pair_c2 6 2 tc_filter_pair_kernel N=1000000 python tools/time_tc.py 10000
cap r2_ncu_pair_125k 6 2 tc_filter_pair_kernel N=125000 python tools/time_tc.py 10000
cap r2_ncu_pair_deep_mode3 6 2 tc_filter_pair_kernel GEN=deep D=96 N=2000000 python tools/time_tc.py 10000
cap r2_ncu_code_filter_c5 2 1 tc_filter_pair_kernel NSPLIT=0 python tools/time_quant.py
echo "== sanitizer"
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_small.py > $O/r2_sanitizer_$tool.log 2>&1; echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_small: done|ok$" $O/r2_sanitizer_$tool.log | tail -8
done

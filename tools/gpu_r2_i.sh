#!/bin/bash
# round-2 GPU call I (1 GPU): ncu --set full captures of the filter kernels, raw pages as CSV
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
cap() { # name skip count -- command (environment assignments via `env` inside the command)
  name=$1; skip=$2; cnt=$3; shift 3
  timeout 900 ncu --set full --clock-control none -k regex:tc_filter_pair_kernel -s $skip -c $cnt -o $O/$name "$@" > $O/$name.log 2>&1; echo "$name rc=$?"
  ncu -i $O/$name.ncu-rep --page raw --csv > $O/${name}_raw.csv 2>/dev/null; ls -la $O/$name.ncu-rep; rm -f $O/$name.ncu-rep
}
cap r2_ncu_pair_c2 6 2 env N=1000000 python tools/time_tc.py 10000
cap r2_ncu_pair_125k 6 2 env N=125000 python tools/time_tc.py 10000
cap r2_ncu_pair_deep_mode3 6 2 env GEN=deep D=96 N=2000000 python tools/time_tc.py 10000
cap r2_ncu_code_filter_c5 2 1 env NSPLIT=0 python tools/time_quant.py

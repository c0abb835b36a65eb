#!/bin/bash
# per-pass times of k_sim_tc (ncu gpu__time_duration) for the knock-out switches: tools/sim_sweep.sh "1 2" "0 8 1 16 25"
for cg in $1; do for dbg in $2; do
  OPB_SIM_CG=$cg OPB_SIM_DBG=$dbg B=64 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:k_sim_tc|k_sim16' -s 4 -c 2 --csv python tools/bench_sim.py 2>/dev/null | grep 'k_sim_tc\|k_sim16' | awk -F'","' -v cg=$cg -v dbg=$dbg '{gsub(/"/,"",$NF); printf "cg=%s dbg=%s %s %s us\n", cg, dbg, substr($5,1,40), $NF}'
done; done

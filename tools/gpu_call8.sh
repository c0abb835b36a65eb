#!/bin/bash
O=gpurun_out
for plan in 4s 3s 2s 2d; do
  echo "== OPB_SIM_P2=$plan"
  for dbg in 0 8 1; do
  OPB_SIM_P2=$plan OPB_SIM_DBG=$dbg B=64 timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:k_sim16' -s 4 -c 2 --csv python tools/bench_sim.py 2>/dev/null | grep 'k_sim16' | awk -F'","' -v dbg=$dbg '{gsub(/"/,"",$NF); printf "dbg=%s %s %s ns\n", dbg, substr($5,1,46), $NF}'
  done
  OPB_SIM_P2=$plan B=64 python tools/bench_sim.py
done > $O/sim_p2.txt 2>&1
cat $O/sim_p2.txt
OPB_SIM_P2=2d timeout 300 python -m pytest tests/test_gpu_matcher.py -x -q -m gpu 2>&1 | tail -3

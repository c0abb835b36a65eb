#!/bin/bash
O=gpurun_out
for dbg in 0 16 4 1 5; do OPB_LIN_DBG=$dbg timeout 120 python tools/bench_linear.py; done > $O/lin_dbg.txt 2>&1
cat $O/lin_dbg.txt

#!/bin/bash
O=gpurun_out
for pf in 0 2 4 8 16; do OPB_LIN_PF=$pf timeout 120 python tools/bench_linear.py; done > $O/lin_pf.txt 2>&1
cat $O/lin_pf.txt
timeout 300 python -m pytest tests/test_gpu_matcher.py tests/test_gpu_linear_tc.py -x -q -m gpu 2>&1 | tail -3

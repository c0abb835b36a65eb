#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_pnp_count|k_pnp_solve' -s 12 -c 4 -o $O/r02e_pnp2 -f \
    python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/ncu_pnp2.log 2>&1

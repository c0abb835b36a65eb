#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
for occ in 2 3 4; do OPB_PNP_SOLVE_OCC=$occ python bench.py --config pnp --no-cpu-baseline > $O/r02f_bench_pnp_occ$occ.json 2> $O/r02f_bench_pnp_occ$occ.err; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r02f_pnp_launches.csv python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/ncu_pnp_l.log 2>&1
python bench.py --no-cpu-baseline > $O/r02f_bench.json 2> $O/r02f_bench.err

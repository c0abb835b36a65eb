#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
python bench.py --config pnp > $O/r02e_bench_pnp.json 2> $O/r02e_bench_pnp.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r02e_pnp_launches.csv python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/ncu_pnp_l.log 2>&1
bash tools/sim_sweep.sh "2" "0 8 1 9 4" > $O/sim_sweep.txt 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_gats_planes -s 5 -c 1 -o $O/r02e_gats -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > $O/ncu_gats.log 2>&1
python bench.py --no-cpu-baseline > $O/r02e_bench.json 2> $O/r02e_bench.err

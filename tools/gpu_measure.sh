#!/bin/bash
# One measurement pass on a B200 box (what profiles/r02d-r02i were produced with): GPU tests, bench lines of the three single-GPU
# configs and the reference arm, the ncu launch list of a step, and ncu --set full captures of the kernels DESIGN.md quotes.
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_measure.sh <tag>'     (outputs under gpurun_out/)
set -x
T=${1:-run}
O=gpurun_out
mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${T}_pytest_gpu.log
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
python bench.py --config large --steps 3 > $O/${T}_bench_large.json 2> $O/${T}_bench_large.err
python bench.py --config pnp > $O/${T}_bench_pnp.json 2> $O/${T}_bench_pnp.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err
python tools/bench_pipeline.py > $O/${T}_pipeline.json 2> $O/${T}_pipeline.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-graph > $O/${T}_ncu_launches.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${T}_pnp_launches.csv \
    python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/${T}_ncu_pnp_launches.log 2>&1
B=64 timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_sim16 -s 4 -c 2 -o $O/${T}_sim16 -f \
    python tools/bench_sim.py > $O/${T}_ncu_sim.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_pnp_solve|k_pnp_count|k_pnp_select_refit' -s 12 -c 5 -o $O/${T}_pnp -f \
    python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/${T}_ncu_pnp.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_gats_planes -s 5 -c 1 -o $O/${T}_gats -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > $O/${T}_ncu_gats.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_linear_tc -s 40 -c 4 -o $O/${T}_linear -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > $O/${T}_ncu_linear.log 2>&1
ls -la $O

#!/bin/bash
# round-2 measurement pass: bench lines of the three single-GPU configs + ncu --set full of the PnP and GATs kernels
set -x
O=gpurun_out
python bench.py > $O/r02d_bench.json 2> $O/r02d_bench.err
python bench.py --config large --steps 3 > $O/r02d_bench_large.json 2> $O/r02d_bench_large.err
python bench.py --config pnp > $O/r02d_bench_pnp.json 2> $O/r02d_bench_pnp.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_pnp_hypotheses|k_pnp_select_refit' -s 6 -c 3 -o $O/r02d_pnp -f \
    python bench.py --config pnp --steps 1 --warmup 1 --no-cpu-baseline > $O/ncu_pnp.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_gats_planes -s 4 -c 1 -o $O/r02d_gats -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > $O/ncu_gats.log 2>&1
ls -la $O

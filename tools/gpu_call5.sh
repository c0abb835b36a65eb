#!/bin/bash
set -x
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
python bench.py > $O/r02g_bench.json 2> $O/r02g_bench.err
python bench.py --config pnp > $O/r02g_bench_pnp.json 2> $O/r02g_bench_pnp.err
python tools/bench_pipeline.py > $O/r02g_pipeline.json 2> $O/r02g_pipeline.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02g_bench_reference.json 2> $O/r02g_bench_reference.err

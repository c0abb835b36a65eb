#!/bin/bash
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_matcher.py -x -q -m gpu 2>&1 | tail -3
for r in 0 1; do OPB_GATS_REG=$r python bench.py --no-cpu-baseline > $O/r02j_bench_gatsreg$r.json 2>/dev/null; done
python - <<'PY'
import json
for r in (0, 1):
    d = json.load(open(f"gpurun_out/r02j_bench_gatsreg{r}.json"))
    print(r, round(d["value"], 1), round(d["ms_per_step"], 3), d["kernels"]["op_gats_leaf_aggregate_planes"])
PY

#!/bin/bash
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q -m gpu 2>&1 | tail -5
for bk in 64 32; do OPB_LIN_BK=$bk timeout 120 python tools/bench_linear.py; done > $O/lin_bk.txt 2>&1
cat $O/lin_bk.txt
timeout 600 python -m pytest tests/test_gpu_matcher.py -x -q -m gpu 2>&1 | tail -3
for bk in 64 32; do OPB_LIN_BK=$bk python bench.py --no-cpu-baseline > $O/r02i_bench_bk$bk.json 2>/dev/null; done
python - <<'PY'
import json
for bk in (64, 32):
    d = json.load(open(f"gpurun_out/r02i_bench_bk{bk}.json"))
    print(bk, round(d["value"], 1), round(d["ms_per_step"], 3), {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items() if k.startswith("op_linear")})
PY

#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > $O/r02i_bench_2gpu.json 2> $O/r02i_bench_2gpu.err
tail -c 600 $O/r02i_bench_2gpu.err
python -c "
import json; d=json.load(open('$O/r02i_bench_2gpu.json')); print(d['value'], d['ms_per_step'], d['ms_per_step_per_rank'], d['e2e']['value'], d['clocks'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 | tail -c 400

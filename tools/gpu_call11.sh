#!/bin/bash
O=gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python tools/bench_pipeline.py > $O/r02i_pipeline.json 2> $O/r02i_pipeline.err; cat $O/r02i_pipeline.json; tail -3 $O/r02i_pipeline.err
python tools/bench_objectdb.py 2>&1 | tail -3

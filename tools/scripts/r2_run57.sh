#!/bin/bash
cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2_t57_full.log
timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_t57_bench.json 2> gpurun_out/r2_t57_bench.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_t57_smoke.log 2>&1

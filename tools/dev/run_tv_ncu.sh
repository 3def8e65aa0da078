#!/bin/bash
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"TvpViolationKernel" --launch-skip 200 -c 3 -o gpurun_out/tv_full_viol -f python tools/bench_configs.py tv1d --n 10000000 --no-solve --no-profile --warmup 3 --steps 70 > gpurun_out/tv_ncu_full.log 2>&1
tail -n 2 gpurun_out/tv_ncu_full.log | cut -c1-200

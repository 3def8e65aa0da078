#!/bin/bash
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"ElemProxVec2Kernel" --launch-skip 30 -c 1 -o gpurun_out/logi_full -f python tools/bench_prox.py > gpurun_out/logi_ncu.log 2>&1
tail -n 2 gpurun_out/logi_ncu.log | cut -c1-300

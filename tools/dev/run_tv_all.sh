#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 6 > gpurun_out/suite.log
cat gpurun_out/suite.log
bash tools/dev/run_tv.sh

#!/bin/bash
# the no-alloc entry point: alone in a fresh process, then at the end of the whole GPU suite
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 15 > gpurun_out/noalloc_alone.log
cat gpurun_out/noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 15 > gpurun_out/noalloc_suite.log
cat gpurun_out/noalloc_suite.log

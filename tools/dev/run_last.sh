#!/bin/bash
timeout 300 python tools/bench_configs.py tv1d --n 10000000 > gpurun_out/tv_1e7_spec.json 2> gpurun_out/tv_1e7_spec.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/tv_1e7_spec.json").read().strip().splitlines()[-1])
print(d["solve"]["num_iterations"], d["solve"]["objective"], d["iterations_per_s"], d["ms_per_iteration"], {k:(v["launches"], round(v["ms_per_iteration"],4)) for k,v in d["classes"].items()})
PY
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4 > gpurun_out/final5_suite.log
cat gpurun_out/final5_suite.log

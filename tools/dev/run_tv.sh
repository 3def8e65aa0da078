#!/bin/bash
# TV-1D solver: parity tests, then the config-1 measurement and its ncu launch list
timeout 600 python -m pytest tests/test_gpu_golden.py tests/test_gpu_solve.py tests/test_gpu_adapter.py tests/test_solvemodule_py3.py -m gpu -x -q -k "tv or TOTAL or total_variation" 2>&1 | tail -n 8 > gpurun_out/tv_pytest.log
cat gpurun_out/tv_pytest.log
EPS_TV_TRACE=1 timeout 300 python tools/bench_configs.py tv1d --n 10000000 > gpurun_out/tv_1e7.json 2> gpurun_out/tv_1e7.err
grep "^\[tv\]" gpurun_out/tv_1e7.err | head -n 30
python - <<'PY'
import json
d=json.loads(open("gpurun_out/tv_1e7.json").read().strip().splitlines()[-1])
print(d["solve"]["num_iterations"], d["solve"]["objective"], d["iterations_per_s"], d["ms_per_iteration"], {k:(v["launches"], round(v["ms_per_iteration"],4)) for k,v in d["classes"].items()})
PY
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/tv_launches.csv python tools/bench_configs.py tv1d --n 10000000 --no-solve --no-profile --warmup 3 --steps 12 > gpurun_out/tv_ncu.log 2>&1
wc -l gpurun_out/tv_launches.csv

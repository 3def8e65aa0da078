#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_golden.py tests/test_gpu_solve.py tests/test_gpu_adapter.py tests/test_solvemodule_py3.py -m gpu -x -q -k "logistic or logreg or LOGISTIC" 2>&1 | tail -n 6 > gpurun_out/logi_pytest.log
cat gpurun_out/logi_pytest.log
timeout 300 python tools/bench_prox.py > gpurun_out/prox_roofline.json 2> gpurun_out/prox_roofline.err
tail -n 2 gpurun_out/prox_roofline.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/prox_roofline.json").read().strip().splitlines()[-1])
for k,v in d["kernels"].items(): print(k, round(v["us_per_launch"],1), round(v["frac_of_measured_hbm_peak"],3))
PY

#!/bin/bash
# end-of-session validation: whole GPU suite, smoke, default bench
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 6 > gpurun_out/final2_suite.log
cat gpurun_out/final2_suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final2_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/final2_smoke.log
timeout 600 python bench.py > gpurun_out/final2_bench.json 2> gpurun_out/final2_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/final2_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d.get("parity_check"), d["cpu_baseline"]["value"], d["clocks"])
PY

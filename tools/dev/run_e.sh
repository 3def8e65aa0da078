python -m pytest tests/test_gpu_golden.py -m gpu -x -q -k "sparse and cg" 2>&1 | tail -3
python bench.py --no-cpu-baseline > gpurun_out/r02e_bench_v2.json 2> gpurun_out/r02e_bench_v2.err
python bench.py --m 125000 --no-e2e --no-cpu-baseline --no-parity-check > gpurun_out/r02e_shard_v2.json 2> /dev/null
EPS_STEP_TRACE=400 python bench.py --m 125000 --no-e2e --no-cpu-baseline --no-parity-check 2>&1 >/dev/null | grep "step trace" > gpurun_out/r02e_shard_trace.txt
EPS_STEP_TRACE=400 python bench.py --no-e2e --no-cpu-baseline --no-parity-check 2>&1 >/dev/null | grep "step trace" > gpurun_out/r02e_full_trace.txt
for f in bench_v2 shard_v2; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02e_$f.json"))
    print("$f", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "cg/step", round(d["config"]["cg_iterations_per_step"],2), "e2e", d["e2e"] and round(d["e2e"]["value"],1), "parity", d.get("parity_check") and d["parity_check"]["ok"], {k:(v["launches"], round(v["ms"],2)) for k,v in d["roofline"]["classes"].items()})
except Exception as e:
    print("$f failed", e)
PY
done
cat gpurun_out/r02e_shard_trace.txt; cat gpurun_out/r02e_full_trace.txt

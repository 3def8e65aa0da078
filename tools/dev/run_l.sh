python -m pytest tests/test_gpu_golden.py tests/test_gpu_maps.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sparse or full" 2>&1 | tail -3
b() { python bench.py --no-e2e --no-cpu-baseline --no-parity-check "$@"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], round(d["value"],1), "it/s  ms/step", round(d["ms_per_step"],3), "cg steps/s", round(d["config"]["cg_steps_per_s"]), {k:(v["launches"], round(1e3*v["ms"]/max(v["launches"],1),1)) for k,v in d["roofline"]["classes"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
b --m 125000 > gpurun_out/r02l_shard_nospan.json 2>/dev/null; show gpurun_out/r02l_shard_nospan.json
EPS_SELL_SPAN=1 b --m 125000 > gpurun_out/r02l_shard_span.json 2>/dev/null; show gpurun_out/r02l_shard_span.json
b --m 250000 > gpurun_out/r02l_shard4_nospan.json 2>/dev/null; show gpurun_out/r02l_shard4_nospan.json
EPS_SELL_SPAN=1 b --m 250000 > gpurun_out/r02l_shard4_span.json 2>/dev/null; show gpurun_out/r02l_shard4_span.json
b > gpurun_out/r02l_full.json 2>/dev/null; show gpurun_out/r02l_full.json
EPS_SELL_SPAN=1 b > gpurun_out/r02l_full_span.json 2>/dev/null; show gpurun_out/r02l_full_span.json
EPS_STEP_TRACE=400 b --m 125000 2>&1 >/dev/null | grep "step trace" | head -8

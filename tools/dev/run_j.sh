python -m pytest tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sparse or full" 2>&1 | tail -3
b() { python bench.py --no-e2e --no-cpu-baseline --no-parity-check "$@"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], round(d["value"],1), "it/s  ms/step", round(d["ms_per_step"],3), "cg steps/s", round(d["config"]["cg_steps_per_s"]), {k:(v["launches"], round(1e3*v["ms"]/max(v["launches"],1),1)) for k,v in d["roofline"]["classes"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
b --m 125000 > gpurun_out/r02j_shard_base.json 2>/dev/null; show gpurun_out/r02j_shard_base.json
for t in 20000 22000 18000; do EPS_SELL_TILE=$t b --m 125000 > gpurun_out/r02j_shard_tile$t.json 2>/dev/null; show gpurun_out/r02j_shard_tile$t.json; done
b > gpurun_out/r02j_full_base.json 2>/dev/null; show gpurun_out/r02j_full_base.json
for t in 20000 22000; do EPS_SELL_TILE=$t b > gpurun_out/r02j_full_tile$t.json 2>/dev/null; show gpurun_out/r02j_full_tile$t.json; done
EPS_STEP_TRACE=400 b --m 125000 2>&1 >/dev/null | grep "step trace" | head -8

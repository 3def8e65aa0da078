# shard-shaped problem on one GPU: staging experiments
b() { python bench.py --m 125000 --no-e2e --no-cpu-baseline --no-parity-check "$@"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], round(d["value"],1), "it/s  ms/step", round(d["ms_per_step"],3), "cg steps/s", round(d["config"]["cg_steps_per_s"]), {k:(v["launches"], round(1e3*v["ms"]/max(v["launches"],1),1)) for k,v in d["roofline"]["classes"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
b > gpurun_out/r02i_base.json 2>/dev/null; show gpurun_out/r02i_base.json
for c in 8192 16384 32768 65536; do EPS_SELL_TMA_CHUNK=$c b > gpurun_out/r02i_chunk$c.json 2>/dev/null; show gpurun_out/r02i_chunk$c.json; done
for t in 12800 16000 20000; do EPS_SELL_TILE=$t b > gpurun_out/r02i_tile$t.json 2>/dev/null; show gpurun_out/r02i_tile$t.json; done
EPS_SELL_TILE=12800 EPS_SELL_TMA_CHUNK=16384 b > gpurun_out/r02i_tile12800_chunk.json 2>/dev/null; show gpurun_out/r02i_tile12800_chunk.json
EPS_STEP_TRACE=400 b 2>&1 >/dev/null | grep "step trace" | head -8
EPS_SELL_TMA_CHUNK=16384 EPS_STEP_TRACE=400 b 2>&1 >/dev/null | grep "step trace" | head -8

b() { python bench.py --no-e2e --no-cpu-baseline --no-parity-check "$@"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], round(d["value"],1), "it/s  ms/step", round(d["ms_per_step"],3), "cg steps/s", round(d["config"]["cg_steps_per_s"]), {k:(v["launches"], round(1e3*v["ms"]/max(v["launches"],1),1)) for k,v in d["roofline"]["classes"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
for t in 12800 14000 16000 18000 20000 25600; do EPS_SELL_TILE=$t b > gpurun_out/r02k_full_tile$t.json 2>/dev/null; show gpurun_out/r02k_full_tile$t.json; done
for t in 12800 14000 16000 20000; do EPS_SELL_TILE=$t b --m 125000 > gpurun_out/r02k_shard_tile$t.json 2>/dev/null; show gpurun_out/r02k_shard_tile$t.json; done

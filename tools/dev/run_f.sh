# 8-GPU: v2 vs v1 and a step trace on rank 0
N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) bench.py --gpus $N "$@"; }
run --steps 20 --warmup 5 > gpurun_out/r02f_n${N}_v2.json 2> gpurun_out/r02f_n${N}_v2.err
EPS_TAIL_V1=1 run --steps 20 --warmup 5 > gpurun_out/r02f_n${N}_v1.json 2> gpurun_out/r02f_n${N}_v1.err
EPS_STEP_TRACE=300 run --steps 20 --warmup 5 --no-e2e 2>&1 >/dev/null | grep "step trace" > gpurun_out/r02f_n${N}_trace.txt
for f in v2 v1; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02f_n${N}_$f.json"))
    print("$f N=$N", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "cg/step", round(d["config"]["cg_iterations_per_step"],2), "e2e", d["e2e"] and round(d["e2e"]["value"],1), "parity", d.get("parity_check") and d["parity_check"]["ok"], {k:(v["launches"], round(v["ms"],2)) for k,v in d["roofline"]["classes"].items()})
except Exception as e:
    print("$f failed", e)
PY
done
cat gpurun_out/r02f_n${N}_trace.txt
tail -3 gpurun_out/r02f_n${N}_v2.err

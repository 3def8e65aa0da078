N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) bench.py --gpus $N "$@"; }
EPS_TIMING=1 run --steps 20 --warmup 5 > gpurun_out/r02m_n${N}.json 2> gpurun_out/r02m_n${N}.err
grep "eps timing" gpurun_out/r02m_n${N}.err | sort | uniq -c | sort -rn | head -12
python - <<PY
import json
d=json.load(open("gpurun_out/r02m_n${N}.json"))
print("N=$N", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "cg/step", round(d["config"]["cg_iterations_per_step"],2), "parity", d.get("parity_check") and d["parity_check"]["ok"])
print({k:(round(v,4) if isinstance(v,float) else v) for k,v in d["e2e"].items() if k in ("value","time_to_tol_s","construct_s","solve_s","init_s","loop_s","iterations")})
PY

N=${1:-4}; M=${2:-500000}; K=${3:-1000}
timeout ${4:-360} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) tools/bench_configs.py group_lasso --rows $M --group 50 --K $K --device-gen --steps 10 --no-profile --reuse-solver > gpurun_out/r02_config_group_lasso_${M}_n$N.json 2> gpurun_out/r02u_gl_n$N.err
python - gpurun_out/r02_config_group_lasso_${M}_n$N.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("N=%d"%d["n_gpus"], round(d["iterations_per_s"],2), "it/s gen %.0fs"%d["data_generation_s"], d.get("solve"), d["solver"][:2])
except Exception as e: print(sys.argv[1], "failed", e)
PY
grep -v "^W1105\|OMP_NUM\|^\*\*\*" gpurun_out/r02u_gl_n$N.err | tail -6 | cut -c1-300

N=${1:-8}
if [ "$N" = "1" ]; then
  tr() { timeout 600 python tools/bench_configs.py "$@"; }
else
  tr() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) tools/bench_configs.py "$@"; }
fi
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1].split('/')[-1], "N=%d"%d["n_gpus"], round(d["iterations_per_s"],2), "it/s gen %.0fs"%d["data_generation_s"], "cg/it", d.get("cg_iterations_per_step"), d.get("solve"), d["solver"][-1], {k:(v["launches"], round(v["us_per_launch"],1), round(v["frac_of_measured_hbm_peak"] or 0,2)) for k,v in d.get("classes",{}).items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
tr logreg --rows 2000000 --cols 200000 --density 1e-3 --steps 20 > gpurun_out/r02_config_logreg_2Mx200k_n$N.json 2> gpurun_out/r02t_logreg_n$N.err; show gpurun_out/r02_config_logreg_2Mx200k_n$N.json; grep -i "error\|Traceback" gpurun_out/r02t_logreg_n$N.err | head -3
tr hinge --rows 1000000 --cols 10000 --device-gen --steps 20 > gpurun_out/r02_config_hinge_1Mx10k_n$N.json 2> gpurun_out/r02t_hinge_n$N.err; show gpurun_out/r02_config_hinge_1Mx10k_n$N.json; grep -i "error\|Traceback" gpurun_out/r02t_hinge_n$N.err | head -3

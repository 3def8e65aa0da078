N=${1:-2}
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) tools/bench_configs.py "$@"; }
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], "N=%d"%d["n_gpus"], round(d["iterations_per_s"],1), "it/s", d.get("solve"), d["solver"][1:], {k:(v["launches"], round(v["us_per_launch"],1), round(v["frac_of_measured_hbm_peak"] or 0,2)) for k,v in d.get("classes",{}).items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
tr logreg --rows 200000 --cols 20000 --density 1e-3 --steps 10 > gpurun_out/r02s_logreg_n$N.json 2> gpurun_out/r02s_logreg_n$N.err; show gpurun_out/r02s_logreg_n$N.json; tail -3 gpurun_out/r02s_logreg_n$N.err | cut -c1-300
python tools/bench_configs.py logreg --m 200000 --n 20000 --density 1e-3 --steps 10 > gpurun_out/r02s_logreg_n1.json 2>/dev/null; show gpurun_out/r02s_logreg_n1.json
tr hinge --rows 100000 --cols 2000 --device-gen --steps 10 > gpurun_out/r02s_hinge_n$N.json 2> gpurun_out/r02s_hinge_n$N.err; show gpurun_out/r02s_hinge_n$N.json; tail -3 gpurun_out/r02s_hinge_n$N.err | cut -c1-300
python tools/bench_configs.py hinge --m 100000 --n 2000 --device-gen --steps 10 > gpurun_out/r02s_hinge_n1.json 2>/dev/null; show gpurun_out/r02s_hinge_n1.json

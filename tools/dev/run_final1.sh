python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r02_pytest_gpu.log; cat gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_n1.json"))
print(round(d["value"],1), d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["time_to_tol_s"], "parity", d["parity_check"]["ok"], d["parity_check"]["max_rel"])
print({k:d["roofline"][k] for k in ("achieved","frac","traffic","ms_per_launch","share_of_step")})
print(d["vs_reference_same_config"]); print(d["clocks"]); print(d["gpu_launches"])
PY
python bench.py --impl reference 2>/dev/null | head -c 400

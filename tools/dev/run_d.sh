python -m pytest tests/test_gpu_golden.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sparse or fullsize or full_size" 2>&1 | tail -15 > gpurun_out/r02d_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r02d_bench_v2.json 2> gpurun_out/r02d_bench_v2.err
EPS_TAIL_V1=1 python bench.py --no-cpu-baseline > gpurun_out/r02d_bench_v1.json 2> gpurun_out/r02d_bench_v1.err
EPS_NO_CARRY=1 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/r02d_bench_v2_nocarry.json 2> /dev/null
python bench.py --m 125000 --no-e2e --no-cpu-baseline --no-parity-check > gpurun_out/r02d_shard_v2.json 2> /dev/null
EPS_TAIL_V1=1 python bench.py --m 125000 --no-e2e --no-cpu-baseline --no-parity-check > gpurun_out/r02d_shard_v1.json 2> /dev/null
tail -5 gpurun_out/r02d_pytest.log
for f in bench_v2 bench_v1 bench_v2_nocarry shard_v2 shard_v1; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02d_$f.json"))
    print("$f", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "cg/step", round(d["config"]["cg_iterations_per_step"],2), "e2e", d["e2e"] and round(d["e2e"]["value"],1), "parity", d.get("parity_check") and d["parity_check"]["ok"], {k:(v["launches"], round(v["ms"],2)) for k,v in d["roofline"]["classes"].items()})
except Exception as e:
    print("$f failed", e)
PY
done

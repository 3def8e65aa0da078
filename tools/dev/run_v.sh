python -m pytest tests/test_gpu_golden.py -m gpu -x -q -k "prox or elementwise or quantile or logistic" 2>&1 | tail -3
python tools/bench_prox.py > gpurun_out/r02_prox_roofline.json 2> gpurun_out/r02v.err; cat gpurun_out/r02_prox_roofline.json
EPS_PROX_SCALAR=1 python tools/bench_prox.py 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:ElemProxVec2 -s 6 -c 2 -o gpurun_out/r02_elemprox_norm1 python tools/bench_prox.py --steps 3 > /dev/null 2> gpurun_out/r02v_ncu.err
ncu -i gpurun_out/r02_elemprox_norm1.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_ElemProxKernel.csv 2>/dev/null
rm -f gpurun_out/r02_elemprox_norm1.ncu-rep
wc -c gpurun_out/r02_ncu_full_ElemProxKernel.csv; tail -3 gpurun_out/r02v_ncu.err

# ncu evidence for round 2 (after the plain bench has exited 0)
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity-check > gpurun_out/r02x_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches_raw.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity-check > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:SellStreamKernel -s 20 -c 2 -o gpurun_out/r02_sell python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity-check > /dev/null 2> gpurun_out/r02x_ncu1.err
ncu -i gpurun_out/r02_sell.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_SellStreamKernel.csv 2>/dev/null
ncu --set full --clock-control none -k regex:CgFusedTail2Kernel -s 10 -c 1 -o gpurun_out/r02_tail python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity-check > /dev/null 2> gpurun_out/r02x_ncu2.err
ncu -i gpurun_out/r02_tail.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_CgFusedTail2Kernel.csv 2>/dev/null
ncu --set full --clock-control none -k regex:ElemProxVec2 -s 1 -c 1 -o gpurun_out/r02_prox python tools/bench_prox.py --steps 2 > /dev/null 2> gpurun_out/r02x_ncu3.err
ncu -i gpurun_out/r02_prox.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_ElemProxVec2Kernel_norm1.csv 2>/dev/null
rm -f gpurun_out/*.ncu-rep
wc -c gpurun_out/r02_ncu_*.csv

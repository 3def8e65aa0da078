N=${1:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 1000)) bench.py --gpus $N "$@"; }
EPS_STEP_TRACE=300 run --steps 20 --warmup 5 --no-e2e 2>&1 >/dev/null | grep "step trace" | sort -s -k4,4n > gpurun_out/r02g_n${N}_trace.txt
cat gpurun_out/r02g_n${N}_trace.txt

# multi-GPU checks on one box: tests for the world sizes the box holds, then bench at each N (both gather forms)
set -x
NG=$(python -c "import torch; print(torch.cuda.device_count())")
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multi.log
for N in ${NS:-2 4 8}; do
  [ $N -le $NG ] || continue
  for G in ${GATHERS:-fused nccl}; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 20 --warmup 5 --gather $G --cpu-sample 0 ${BENCH_ARGS:-} > gpurun_out/bench_n${N}_$G.json 2> gpurun_out/bench_n${N}_$G.err
    echo "bench N=$N $G rc=$?"
    python -c "import json; d=json.loads(open('gpurun_out/bench_n${N}_$G.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['ms_per_step'], d['value']/1e9, d['roofline']['kernel_ms'], d['roofline']['after_kernel_ms'], d['verified']['indices'], d['verified']['counts'], d['roofline'].get('fused_gather'), d['e2e']['value'] if d['e2e'] else None)" || tail -c 1500 gpurun_out/bench_n${N}_$G.err
  done
done

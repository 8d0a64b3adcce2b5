# ncu --set full capture of one classify launch on the bench workload (after the plain run has exited 0)
set -x
NAME=${1:-r02_classify}
B="python bench.py --steps 2 --warmup 3 --e2e-steps 0 --cpu-sample 0 --json-nodes 0 --seam-nodes 0"
$B > gpurun_out/${NAME}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 3 -c 1 -f -o gpurun_out/$NAME $B > gpurun_out/${NAME}_ncu.log 2>&1
cat gpurun_out/${NAME}_plain.log | tail -1 | cut -c1-400

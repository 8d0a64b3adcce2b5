# Round-2 ncu evidence (one GPU).  Each ncu run follows a plain run of the same command line that exited 0.
#   1. launch list of the default bench (kernel shares of a step)
#   2. --set full capture of one classify launch on the bench workload
#   3. --set full captures of the three large kernels of the JSON scan
set -x
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --cpu-sample 0 --json-nodes 1000000 --json-steps 2 --json-cpu-sample 0 --seam-nodes 0"
$B > gpurun_out/r02_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_n1.csv $B > gpurun_out/r02_ncu_launches.log 2>&1
$B > gpurun_out/r02_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 4 -c 1 -f -o gpurun_out/r02_classify_10m $B > gpurun_out/r02_ncu_classify.log 2>&1
$B > gpurun_out/r02_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:index_kernel|measure_rows_kernel|write_rows_kernel' -s 9 -c 3 -f -o gpurun_out/r02_json_1m $B > gpurun_out/r02_ncu_json.log 2>&1
ls -la gpurun_out | grep r02_

set -x
B="python bench.py --steps 2 --warmup 3 --e2e-steps 0 --cpu-sample 0 --json-nodes 1000000 --json-steps 1 --json-cpu-sample 0 --seam-nodes 0"
$B > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 3 -c 1 -f -o gpurun_out/r02_classify_base $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/prof_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:index_kernel|measure_rows_kernel|write_rows_kernel' -s 9 -c 3 -f -o gpurun_out/r02_json_base $B > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu1.log gpurun_out/ncu2.log
ls -la gpurun_out

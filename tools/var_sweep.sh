# sweep kernel variants on the bench workload; every command under its own timeout
N=${N:-10000000}
for v in ${VARIANTS:-t224 t288g2 t256g2}; do
  echo "== $v"
  GNC_VARIANT=$v timeout 90 python bench.py --nodes $N --json-nodes 0 --seam-nodes 0 --cpu-sample 0 --e2e-steps 0 > gpurun_out/sweep_$v.json 2> gpurun_out/sweep_$v.err
  echo "rc=$?"; tail -c 600 gpurun_out/sweep_$v.err
  python -c "import sys,json; d=json.loads(open('gpurun_out/sweep_$v.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['roofline']['frac'], d['roofline']['kernel_ms'], d['verified']['indices'], d['verified']['counts'])" 2>/dev/null
done

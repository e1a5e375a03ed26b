mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 100 --warmup 10 --cpu-seconds 1 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print("value %.3f M/s  ms/step %.4f  e2e %.3f M/s  step frac %.4f" % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6, d['roofline_step']['frac']))
for k,v in d['kernels'].items(): print("  %-22s %8.2f us  %s GB/s" % (k, v['avg_us'], v['GBps']))
PY

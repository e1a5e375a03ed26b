mkdir -p gpurun_out
SMIL_FWD_ASSIGN=anchor python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu 2>&1 | tail -2
bash tools_gpu/run_fwd_assign.sh uniform anchor
for A in uniform anchor uniform anchor; do
SMIL_FWD_ASSIGN=$A python bench.py --steps 100 --warmup 10 --cpu-seconds 0.3 > gpurun_out/bench_$A.log 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$A.log').read().strip().splitlines()[-1]); k=d['kernels']
print("$A: value %.3f M/s step %.4f ms skin_fwd %.2f us" % (d['value']/1e6, d['ms_per_step'], k['skin_fwd_kernel']['avg_us']))
PY
done

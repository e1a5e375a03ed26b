mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
bash tools_gpu/run_fwd_assign.sh uniform anchor
for A in uniform anchor uniform anchor; do
SMIL_FWD_ASSIGN=$A python bench.py --steps 100 --warmup 10 --cpu-seconds 0.3 > gpurun_out/bench_$A.log 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$A.log').read().strip().splitlines()[-1]); k=d['kernels']
print("$A: value %.3f M/s step %.4f ms e2e %.3f skin_fwd %.2f us" % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6, k['skin_fwd_kernel']['avg_us']))
PY
done
SMIL_FWD_ASSIGN=anchor python bench.py --rig mouse --steps 50 --warmup 5 --cpu-seconds 0.3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mouse anchor: value %.3f M/s skin_fwd %.2f us'%(d['value']/1e6, d['kernels']['skin_fwd_kernel']['avg_us']))"

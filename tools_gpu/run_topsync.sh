mkdir -p gpurun_out
for rep in 1 2; do for T in 0 1; do for R in stick mouse; do
SMIL_FWD_TOPSYNC=$T python bench.py --rig $R --steps 100 --warmup 10 --cpu-seconds 0.2 > gpurun_out/b_ts.log 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/b_ts.log').read().strip().splitlines()[-1]); k=d['kernels']
print("topsync $T $R: value %.3f M/s step %.4f ms skin_fwd %.2f us" % (d['value']/1e6, d['ms_per_step'], k['skin_fwd_kernel']['avg_us']))
PY
done; done; done

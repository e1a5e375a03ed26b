# what each section of the pose kernels costs: the library must be built with SMIL_WHATIF=1 (results are wrong when a bit is set)
mkdir -p gpurun_out
for W in 0 1 2 4 8 16 256 512 1024 2048 4096 8192; do
  SMIL_POSE_WHATIF=$W python bench.py --steps 50 --warmup 5 --cpu-seconds 0.3 > gpurun_out/whatif_$W.log 2> gpurun_out/whatif_$W.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/whatif_$W.log').read().strip().splitlines()[-1])
k=d['kernels']
print("whatif %5d: step %.4f ms  pose_fwd %6.2f  skin_fwd %6.2f  skin_bwd %6.2f  pose_bwd %6.2f" % ($W, d['ms_per_step'], k['pose_fwd_kernel']['avg_us'], k['skin_fwd_kernel']['avg_us'], k['skin_bwd_kernel']['avg_us'], [v for n,v in k.items() if n.startswith('pose_bwd')][0]['avg_us']))
PY
done

mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c5_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 default
  SMIL_POSE_V2=0 python tools_gpu/kbench.py stick 4096 pose_v1
  python tools_gpu/kbench.py mouse 4096 mouse_default ) > gpurun_out/c5_kbench.jsonl 2> gpurun_out/c5_kbench.err
cat gpurun_out/c5_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c5_kbench.err | tail -3
python tools_quick_bench.py stick 4096 > gpurun_out/c5_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pose_ -s 8 -c 2 -f -o gpurun_out/r02_pose_a python tools_quick_bench.py stick 4096 > gpurun_out/c5_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/c5_ncu.log

mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/c4_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/c4_pytest.log
( python tools_gpu/kbench.py stick 4096 default
  SMIL_POSE_V2=0 python tools_gpu/kbench.py stick 4096 pose_v1
  SMIL_BWD_VM=0 python tools_gpu/kbench.py stick 4096 item_major
  SMIL_VM_CHUNK=252 python tools_gpu/kbench.py stick 4096 vm_chunk252
  SMIL_VM_CHUNK=192 python tools_gpu/kbench.py stick 4096 vm_chunk192
  python tools_gpu/kbench.py mouse 4096 mouse_default
  SMIL_POSE_V2=0 SMIL_BWD_VM=0 python tools_gpu/kbench.py mouse 4096 mouse_old ) > gpurun_out/c4_kbench.jsonl 2> gpurun_out/c4_kbench.err
cat gpurun_out/c4_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c4_kbench.err | tail -3

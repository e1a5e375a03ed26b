# round 2, call 2: vertex-major backward skin pass -- parity + first timings, against the item-major pass
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_losses.py -x -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c2_pytest.log
( SMIL_BWD_VM=0 python tools_gpu/kbench.py stick 4096 item_major
  python tools_gpu/kbench.py stick 4096 vm_default
  SMIL_BWD_STAGES=2 python tools_gpu/kbench.py stick 4096 vm_stages2
  SMIL_BWD_CPS=3 python tools_gpu/kbench.py stick 4096 vm_cps3
  SMIL_VM_CHUNK=192 python tools_gpu/kbench.py stick 4096 vm_chunk192
  SMIL_VM_CHUNK=128 python tools_gpu/kbench.py stick 4096 vm_chunk128
  SMIL_VM_CHUNK=252 python tools_gpu/kbench.py stick 4096 vm_chunk252
  SMIL_VM_CUT_TOL=0.3 python tools_gpu/kbench.py stick 4096 vm_tol03
  SMIL_BWD_VM=0 python tools_gpu/kbench.py mouse 4096 mouse_item_major
  python tools_gpu/kbench.py mouse 4096 mouse_vm ) > gpurun_out/c2_kbench.jsonl 2> gpurun_out/c2_kbench.err
cat gpurun_out/c2_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
tail -3 gpurun_out/c2_kbench.err

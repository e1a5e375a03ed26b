mkdir -p gpurun_out
python -m pytest tests/test_gpu_losses.py tests/test_gpu_parity.py -q -k "not variants" > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/c9_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 nw4_c168
  for c in 36 44 52 64 84; do SMIL_VM_WARPS=1 SMIL_VM_CHUNK=$c python tools_gpu/kbench.py stick 4096 nw1_c$c; done
  for c in 64 84 104 128; do SMIL_VM_WARPS=2 SMIL_VM_CHUNK=$c python tools_gpu/kbench.py stick 4096 nw2_c$c; done
  SMIL_VM_WARPS=1 SMIL_VM_CHUNK=44 python tools_gpu/kbench.py mouse 4096 mouse_nw1_c44
  SMIL_VM_WARPS=2 SMIL_VM_CHUNK=84 python tools_gpu/kbench.py mouse 4096 mouse_nw2_c84 ) > gpurun_out/c9_kbench.jsonl 2> gpurun_out/c9_kbench.err
cat gpurun_out/c9_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c9_kbench.err | tail -3

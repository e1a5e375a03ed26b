mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/c8_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/c8_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 default_atomic_ph1
  SMIL_VM_PHASES=2 python tools_gpu/kbench.py stick 4096 atomic_ph2
  SMIL_VM_PHASES=3 python tools_gpu/kbench.py stick 4096 atomic_ph3
  SMIL_VM_PHASES=2 SMIL_VM_CHUNK=128 python tools_gpu/kbench.py stick 4096 atomic_ph2_chunk128
  SMIL_VM_PHASES=2 SMIL_VM_CHUNK=200 python tools_gpu/kbench.py stick 4096 atomic_ph2_chunk200
  SMIL_BWD_ATOMIC=0 python tools_gpu/kbench.py stick 4096 slots_ph1
  python tools_gpu/kbench.py mouse 4096 mouse_atomic_ph1
  SMIL_VM_PHASES=2 python tools_gpu/kbench.py mouse 4096 mouse_atomic_ph2 ) > gpurun_out/c8_kbench.jsonl 2> gpurun_out/c8_kbench.err
cat gpurun_out/c8_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c8_kbench.err | tail -3

mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/c6_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 default
  for ns in 1000 2000 3000 4000 6000; do SMIL_BWD_STAGGER_NS=$ns python tools_gpu/kbench.py stick 4096 stagger_$ns; done
  SMIL_BWD_STAGGER_NS=3000 SMIL_VM_CHUNK=252 python tools_gpu/kbench.py stick 4096 stagger_3000_chunk252
  SMIL_BWD_STAGGER_NS=3000 SMIL_BWD_VM=0 python tools_gpu/kbench.py stick 4096 item_major
  SMIL_BWD_STAGGER_NS=6000 python tools_gpu/kbench.py mouse 4096 mouse_stagger_6000
  SMIL_BWD_STAGGER_NS=12000 python tools_gpu/kbench.py mouse 4096 mouse_stagger_12000 ) > gpurun_out/c6_kbench.jsonl 2> gpurun_out/c6_kbench.err
cat gpurun_out/c6_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c6_kbench.err | tail -3

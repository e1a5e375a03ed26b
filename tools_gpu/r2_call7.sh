mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_pipeline.py tests/test_gpu_dropin.py -q > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/c7_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 default
  SMIL_BWD_SLOT_REDUCE=0 python tools_gpu/kbench.py stick 4096 no_slot_reduce
  SMIL_VM_CHUNK=252 python tools_gpu/kbench.py stick 4096 chunk252
  SMIL_VM_PHASES=1 SMIL_VM_CHUNK=336 python tools_gpu/kbench.py stick 4096 chunk336
  python tools_gpu/kbench.py mouse 4096 mouse_default ) > gpurun_out/c7_kbench.jsonl 2> gpurun_out/c7_kbench.err
cat gpurun_out/c7_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c7_kbench.err | tail -3
python tools_quick_bench.py stick 4096 > gpurun_out/c7_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:skin_bwd_vm -s 4 -c 1 -f -o gpurun_out/r02_vm_b python tools_quick_bench.py stick 4096 > gpurun_out/c7_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/c7_ncu.log

mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/c10_pytest.log | cut -c1-300
( python tools_gpu/kbench.py stick 4096 default
  SMIL_FWD_SINGLE_FIRST=0 python tools_gpu/kbench.py stick 4096 fwd_single_first_off
  SMIL_FWD_GP=1 python tools_gpu/kbench.py stick 4096 fwd_gp1
  SMIL_FWD_GP=4 python tools_gpu/kbench.py stick 4096 fwd_gp4
  SMIL_FWD_CPS=2 python tools_gpu/kbench.py stick 4096 fwd_cps2
  python tools_gpu/kbench.py mouse 4096 mouse_default
  SMIL_FWD_SINGLE_FIRST=0 python tools_gpu/kbench.py mouse 4096 mouse_single_first_off ) > gpurun_out/c10_kbench.jsonl 2> gpurun_out/c10_kbench.err
cat gpurun_out/c10_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
grep -v Warning gpurun_out/c10_kbench.err | tail -3

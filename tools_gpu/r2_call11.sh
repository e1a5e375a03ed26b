mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/c11_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/c11_pytest.log | cut -c1-300
python __graft_entry__.py --smoke 2>&1 | tail -2
( python tools_gpu/kbench.py stick 4096 default
  python tools_gpu/kbench.py mouse 4096 mouse_default ) > gpurun_out/c11_kbench.jsonl 2> gpurun_out/c11_kbench.err
cat gpurun_out/c11_kbench.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d=json.loads(l); print(d['tag'], d['ms_step'], d['Mmesh_s'], d['kernels_us'])"
python bench.py --steps 100 --warmup 10 > gpurun_out/c11_bench.log 2> gpurun_out/c11_bench.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/c11_bench.log; tail -3 gpurun_out/c11_bench.err

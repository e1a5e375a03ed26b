mkdir -p gpurun_out
export SMIL_FWD_GP=2 SMIL_FWD_CPS=3
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for st in 1 2; do
  export SMIL_BWD_STAGES=$st
  python bench.py --steps 50 --warmup 5 --cpu-seconds 0.5 > gpurun_out/bench_s$st.log 2>/dev/null
  echo "STAGES=$st"; python tools_gpu/show_bench.py gpurun_out/bench_s$st.log
done
SMIL_BWD_STAGES=1 python -m pytest tests -m gpu -x -q 2>&1 | tail -2

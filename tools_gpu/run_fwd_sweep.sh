mkdir -p gpurun_out
for d in 1; do for gp in 1 2 4; do for cps in 2 3 4 6 8; do
  echo "DIRECT=$d GP=$gp CPS=$cps: $(SMIL_FWD_DIRECT=$d SMIL_FWD_GP=$gp SMIL_FWD_CPS=$cps python tools_quick_bench.py stick 4096 2>&1 | head -1)"
done; done; done
SMIL_FWD_DIRECT=1 python -m pytest tests -m gpu -x -q 2>&1 | tail -2

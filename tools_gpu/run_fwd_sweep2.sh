# forward pass tuning knobs re-checked under the v6 thread assignment (fwd-only time of the quick bench)
for gp in 1 2 4 8; do for cps in 2 3 4 5; do
  echo "GP=$gp CPS=$cps: $(SMIL_FWD_GP=$gp SMIL_FWD_CPS=$cps python tools_quick_bench.py stick 4096 2>&1 | head -1)"
done; done

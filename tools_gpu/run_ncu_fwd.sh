mkdir -p gpurun_out
export SMIL_FWD_GP=2 SMIL_FWD_CPS=3
python tools_quick_bench.py stick 4096 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:skin_fwd -s 10 -c 1 -f -o gpurun_out/prof_fwd_tma python tools_quick_bench.py stick 4096 > gpurun_out/ncu_a.log 2>&1
echo "rc=$?"
SMIL_FWD_DIRECT=1 ncu --set full --clock-control none --import-source on -k regex:skin_fwd -s 10 -c 1 -f -o gpurun_out/prof_fwd_direct python tools_quick_bench.py stick 4096 > gpurun_out/ncu_b.log 2>&1
echo "rc=$?"
cat gpurun_out/plain.log

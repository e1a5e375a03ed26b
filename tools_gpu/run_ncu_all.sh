# one ncu --set full capture of the four kernels of a fused step (quick bench, eager launches)
mkdir -p gpurun_out
python tools_quick_bench.py stick 4096 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -s 202 -c 4 -f -o gpurun_out/${1:-prof_all} python tools_quick_bench.py stick 4096 > gpurun_out/ncu_all.log 2>&1
echo "rc=$?"; cat gpurun_out/plain.log; tail -2 gpurun_out/ncu_all.log

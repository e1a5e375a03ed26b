mkdir -p gpurun_out
set -x
python tools_quick_bench.py stick 4096 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -s 100 -c 8 -f -o gpurun_out/prof_v1 python tools_quick_bench.py stick 4096 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"
cat gpurun_out/plain.log
tail -3 gpurun_out/ncu_full.log

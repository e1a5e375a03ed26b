# usage: run_ncu_k.sh <kernel regex> <out name>
mkdir -p gpurun_out
python tools_quick_bench.py stick 4096 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$1 -s 3 -c 1 -f -o gpurun_out/$2 python tools_quick_bench.py stick 4096 > gpurun_out/ncu_k.log 2>&1
echo "rc=$?"; cat gpurun_out/plain.log; tail -2 gpurun_out/ncu_k.log

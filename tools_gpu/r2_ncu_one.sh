# one ncu --set full capture of a kernel of the fused step: usage r2_ncu_one.sh <kernel regex> <out name> [rig]
mkdir -p gpurun_out
K=${1:-skin_bwd}; OUT=${2:-prof}; RIG=${3:-stick}
python tools_quick_bench.py $RIG 4096 > gpurun_out/${OUT}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 1 -f -o gpurun_out/$OUT python tools_quick_bench.py $RIG 4096 > gpurun_out/${OUT}_ncu.log 2>&1
echo "rc=$?"; cat gpurun_out/${OUT}_plain.log | tail -3; tail -2 gpurun_out/${OUT}_ncu.log

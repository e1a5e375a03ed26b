# configs[3] on N GPUs (N = $1, default 2): frames sharded, one NCCL all-reduce per step; then the weak-scaling bench line
N=${1:-2}
mkdir -p gpurun_out
python -m pytest tests/test_fitter_step.py -x -q -m gpu 2>&1 | tail -3
for G in 0 1; do
python tools_gpu/fitter_multigpu.py --frames 65536 --steps 10 --warmup 3 --graph $G > gpurun_out/fitter_1gpu_g$G.json 2> gpurun_out/fitter_1gpu.err; echo "1gpu graph=$G rc=$?"; tail -1 gpurun_out/fitter_1gpu_g$G.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools_gpu/fitter_multigpu.py --frames 65536 --steps 10 --warmup 3 --graph $G > gpurun_out/fitter_${N}gpu_g$G.json 2> gpurun_out/fitter_${N}gpu.err; echo "${N}gpu graph=$G rc=$?"; tail -1 gpurun_out/fitter_${N}gpu_g$G.json; tail -3 gpurun_out/fitter_${N}gpu.err
done
if [ "$2" = "bench" ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 100 --warmup 10 --cpu-seconds 1 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_${N}gpu.json | cut -c1-600
fi

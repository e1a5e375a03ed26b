# configs[3] on N GPUs (N = $1, default 2): frames sharded, one NCCL all-reduce per step; then the weak-scaling bench line
N=${1:-2}
mkdir -p gpurun_out
python tools_gpu/fitter_multigpu.py --frames 65536 --steps 10 --warmup 3 > gpurun_out/fitter_1gpu.json 2> gpurun_out/fitter_1gpu.err; echo "1gpu rc=$?"; tail -1 gpurun_out/fitter_1gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools_gpu/fitter_multigpu.py --frames 65536 --steps 10 --warmup 3 > gpurun_out/fitter_${N}gpu.json 2> gpurun_out/fitter_${N}gpu.err; echo "${N}gpu rc=$?"; tail -1 gpurun_out/fitter_${N}gpu.json; tail -3 gpurun_out/fitter_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 100 --warmup 10 --cpu-seconds 1 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"; tail -1 gpurun_out/bench_${N}gpu.json | cut -c1-600

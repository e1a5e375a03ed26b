# round 2, call 1: FMA-pipe microbenchmarks + the round-1 build re-measured on this pool (baseline for round 2)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt 2>&1
./tools_micro/issue_bench > gpurun_out/c1_issue_bench.txt 2>&1
./tools_micro/fma_bench > gpurun_out/c1_fma_bench.txt 2>&1
python bench.py --steps 100 --warmup 10 --cpu-seconds 1 > gpurun_out/c1_bench.log 2> gpurun_out/c1_bench.err; echo "bench rc=$?"
python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c1_pytest_gpu.log
cat gpurun_out/c1_issue_bench.txt

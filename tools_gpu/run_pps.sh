mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python bench.py --steps 100 --warmup 10 --cpu-seconds 0.3 > gpurun_out/bench_pps.log 2>gpurun_out/bench_pps.err; tail -2 gpurun_out/bench_pps.err
python tools_gpu/show_bench.py gpurun_out/bench_pps.log
python bench.py --rig mouse --steps 50 --warmup 5 --cpu-seconds 0.3 > gpurun_out/bench_pps_mouse.log 2>/dev/null; python tools_gpu/show_bench.py gpurun_out/bench_pps_mouse.log

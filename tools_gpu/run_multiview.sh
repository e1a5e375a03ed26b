# configs[2] (mouse, 18 cameras, 8192 frames) and configs[4] (STICK, 5 cameras, batch 1024, PCA scale/translation) timings
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
python tools_gpu/multiview_bench.py --rig mouse --static 1 --views 18 --frames 8192 > gpurun_out/mv_mouse_1gpu.json 2> gpurun_out/mv_mouse_1gpu.err; echo "mouse 8192 rc=$?"; tail -1 gpurun_out/mv_mouse_1gpu.json; tail -2 gpurun_out/mv_mouse_1gpu.err
python tools_gpu/multiview_bench.py --rig mouse --static 1 --views 18 --frames 1024 > gpurun_out/mv_mouse_1gpu_1024.json 2> gpurun_out/mv_mouse_1gpu_1024.err; echo "mouse 1024 rc=$?"; tail -1 gpurun_out/mv_mouse_1gpu_1024.json
python tools_gpu/multiview_bench.py --rig stick --static 0 --views 5 --frames 1024 --pca 1 --propagate 1 > gpurun_out/mv_stick_1gpu.json 2> gpurun_out/mv_stick_1gpu.err; echo "stick rc=$?"; tail -1 gpurun_out/mv_stick_1gpu.json; tail -2 gpurun_out/mv_stick_1gpu.err
python bench.py --rig mouse --steps 50 --warmup 5 --cpu-seconds 1 > gpurun_out/bench_mouse.json 2> gpurun_out/bench_mouse.err; echo "bench mouse rc=$?"; tail -1 gpurun_out/bench_mouse.json | cut -c1-1500
else
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools_gpu/multiview_bench.py --rig mouse --static 1 --views 18 --frames 8192 > gpurun_out/mv_mouse_${N}gpu.json 2> gpurun_out/mv_mouse_${N}gpu.err; echo "mouse ${N}gpu rc=$?"; tail -1 gpurun_out/mv_mouse_${N}gpu.json; tail -2 gpurun_out/mv_mouse_${N}gpu.err
fi

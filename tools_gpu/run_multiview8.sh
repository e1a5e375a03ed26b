mkdir -p gpurun_out
python tools_gpu/multiview_bench.py --rig stick --static 0 --views 5 --frames 1024 --pca 1 --propagate 1 > gpurun_out/mv_stick_1gpu.json 2> gpurun_out/mv_stick_1gpu.err; echo "stick rc=$?"; tail -1 gpurun_out/mv_stick_1gpu.json; tail -2 gpurun_out/mv_stick_1gpu.err
python -m pytest tests/test_gpu_fullsize.py tests/test_fitter_step.py -x -q -m gpu 2>&1 | tail -2
bash tools_gpu/run_multiview.sh 8

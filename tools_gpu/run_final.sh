mkdir -p gpurun_out
bash tools_gpu/run1.sh
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 | cut -c1-700
bash tools_gpu/run_ncu_all.sh prof_v7 | tail -4

# usage: run_chunk_sweep.sh <chunk> ...   parity tests + bench per backward tile size
mkdir -p gpurun_out
for c in "$@"; do
  echo "== SMIL_BWD_CHUNK=$c"
  SMIL_BWD_CHUNK=$c python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -1
  SMIL_BWD_CHUNK=$c python bench.py --steps 50 --warmup 5 --cpu-seconds 0.5 > gpurun_out/bench_chunk_$c.log 2>/dev/null
  python tools_gpu/show_bench.py gpurun_out/bench_chunk_$c.log | grep "value\|skin_bwd"
done

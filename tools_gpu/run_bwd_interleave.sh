mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -m gpu 2>&1 | tail -2
M=l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum,gpu__time_duration.sum
for I in 0 1; do
  export SMIL_BWD_INTERLEAVE=$I
  ncu --metrics $M --clock-control none -k regex:skin_bwd -s 10 -c 1 --csv --log-file gpurun_out/bwd_il_$I.csv python tools_quick_bench.py stick 4096 > /dev/null 2>&1
  echo "== interleave $I"; python - <<PY
import csv
for r in csv.reader(open("gpurun_out/bwd_il_$I.csv")):
    if len(r) > 5 and r[0].isdigit(): print("  ", r[-3], r[-1])
PY
done
for I in 0 1 0 1; do
SMIL_BWD_INTERLEAVE=$I python bench.py --steps 100 --warmup 10 --cpu-seconds 0.3 > gpurun_out/bench_il$I.log 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_il$I.log').read().strip().splitlines()[-1]); k=d['kernels']
print("interleave $I: value %.3f M/s step %.4f ms skin_bwd %.2f us" % (d['value']/1e6, d['ms_per_step'], k['skin_bwd_kernel']['avg_us']))
PY
done

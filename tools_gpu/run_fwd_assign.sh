# shared-memory wavefront split (loads / stores) and duration of skin_fwd for the thread-assignment variants
mkdir -p gpurun_out
M=l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,gpu__time_duration.sum
for A in ${@:-uniform spread}; do
  export SMIL_FWD_ASSIGN=$A
  python tools_quick_bench.py stick 4096 > gpurun_out/plain_$A.log 2>&1
  ncu --metrics $M --clock-control none -k regex:skin_fwd -s 10 -c 1 --csv --log-file gpurun_out/fwd_assign_$A.csv python tools_quick_bench.py stick 4096 > /dev/null 2>&1
  echo "== $A"; cat gpurun_out/plain_$A.log; python - <<PY
import csv
for r in csv.reader(open("gpurun_out/fwd_assign_$A.csv")):
    if len(r) > 5 and r[0].isdigit(): print("  ", r[-3], r[-1])
PY
done

mkdir -p gpurun_out
./tools_micro/lds_gather > gpurun_out/lds_names.txt
ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum --csv --log-file gpurun_out/lds.csv ./tools_micro/lds_gather > /dev/null 2>&1
python - <<'PY'
import csv
names=[l.strip() for l in open("gpurun_out/lds_names.txt") if l.startswith("launch")]
vals={}
for r in csv.reader(open("gpurun_out/lds.csv")):
    if len(r)>5 and r[0].isdigit(): vals.setdefault(int(r[0]),{})[r[-3]]=float(r[-1].replace(",",""))
for i,n in enumerate(names):
    v=vals.get(i,{})
    wf=v.get("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",0); ins=v.get("smsp__inst_executed_op_shared_ld.sum",1)
    print("%-60s wavefronts/load %.2f"%(n, wf/max(ins,1)))
PY

mkdir -p gpurun_out
python bench.py --steps 50 --warmup 5 --cpu-seconds 1 --no-torch-gpu > gpurun_out/s8_n1.json 2> gpurun_out/s8_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 50 --warmup 5 --cpu-seconds 1 --no-torch-gpu > gpurun_out/s8_n$N.json 2> gpurun_out/s8_n$N.err; echo "n$N rc=$?"
done
python - <<'PY'
import json
base=None
for N in (1,2,4,8):
    try:
        d=json.loads(open("gpurun_out/s8_n%d.json"%N).read().strip().splitlines()[-1])
    except Exception as e:
        print(N,"failed",e); continue
    if N==1: base=d
    f=d.get("fitter") or {}
    ff=d.get("fitter_full") or {}
    print(N, "value %.2fM eff %.3f | single %.2fM | e2e %.2fM eff %.3f (%.4f ms) | fitter %s full %s"%(d["value"]/1e6, d["value"]/base["value"]/N, d["value_single_stream"]["value"]/1e6, d["e2e"]["value"]/1e6, d["e2e"]["value"]/base["e2e"]["value"]/N, d["e2e"]["ms_per_step"], f.get("value"), ff.get("value")))
PY
nvidia-smi topo -m | head -12; lscpu | grep -i "numa\|^CPU(s)"

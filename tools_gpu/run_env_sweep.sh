# usage: run_env_sweep.sh "VAR=a VAR2=b" "VAR=c" ...   (one bench per quoted env set)
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs python bench.py --steps 50 --warmup 5 --cpu-seconds 0.5 > gpurun_out/bench_sweep_$i.log 2>/dev/null
  echo "== $envs"; python tools_gpu/show_bench.py gpurun_out/bench_sweep_$i.log
done

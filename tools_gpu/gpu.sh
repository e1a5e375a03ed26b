#!/usr/bin/env bash
# The GPU-box recipes of this repo in one place; run as `gpurun [--gpus N] -- 'bash tools_gpu/gpu.sh <cmd> [args]'`.
#   tests                         pytest -m gpu, smoke(), then the N=1 bench line
#   bench [bench.py args]         N=1 bench line -> gpurun_out/bench.json
#   multi N [bench.py args]       torchrun bench at N ranks + the reference arm -> gpurun_out/bench_nN.json, ref_nN.json
#   kbench RIG B TAG [ENV=V ...]  graph-replay step time + per-kernel event times for one tuning set (tools_gpu/kbench.py)
#   sweep "ENV=a ENV2=b" ...      one bench line per quoted SMIL_* set
#   ncu-list                      launch list of the bench command (gpu__time_duration.sum) -> gpurun_out/launches.csv
#   ncu-full KERNEL_REGEX OUT [RIG]   one `--set full` capture of one kernel of the eager step -> gpurun_out/OUT.ncu-rep
#   ncu-step OUT                  one `--set full` capture of the four kernels of a step
#   guards                        the guard-band / poisoned-scratch / bit-identical-relaunch tests (compute-sanitizer is
#                                 closed on this pool: profiles/r02_sanitizer.md)
# ncu runs only after the same command exited 0 without it; a number printed under ncu is never a bench value.
set -u
mkdir -p gpurun_out
cmd=${1:-tests}; shift || true
case "$cmd" in
  tests)
    python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-300
    python __graft_entry__.py --smoke 2>&1 | tail -2
    python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench.json; tail -3 gpurun_out/bench.err ;;
  bench)
    python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench.json; tail -3 gpurun_out/bench.err ;;
  multi)
    N=$1; shift
    run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
    run 29517 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "n$N rc=$?"; tail -c 3500 gpurun_out/bench_n$N.json; grep diag gpurun_out/bench_n$N.err
    run 29519 --impl reference --steps 3 --warmup 1 > gpurun_out/ref_n$N.json 2> gpurun_out/ref_n$N.err; echo "ref rc=$?"; tail -c 600 gpurun_out/ref_n$N.json ;;
  kbench)
    RIG=$1; B=$2; TAG=$3; shift 3
    env "$@" python tools_gpu/kbench.py $RIG $B $TAG | tee -a gpurun_out/kbench.jsonl ;;
  sweep)
    i=0
    for envs in "$@"; do
      i=$((i+1))
      env $envs python bench.py --steps 50 --warmup 5 --cpu-seconds 0.5 --no-torch-gpu > gpurun_out/bench_sweep_$i.json 2>/dev/null
      echo "== $envs"; python tools_gpu/show_bench.py gpurun_out/bench_sweep_$i.json
    done ;;
  ncu-list)
    python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 --no-torch-gpu > gpurun_out/ncu_list_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
    ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
        python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 --no-torch-gpu > gpurun_out/ncu_list.log 2>&1
    echo "rc=$?"; grep -c smil gpurun_out/launches.csv ;;
  ncu-full)
    K=${1:-skin_bwd}; OUT=${2:-prof}; RIG=${3:-stick}
    python tools/quick_bench.py $RIG 4096 > gpurun_out/${OUT}_plain.log 2>&1 && \
    ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 1 -f -o gpurun_out/$OUT python tools/quick_bench.py $RIG 4096 > gpurun_out/${OUT}_ncu.log 2>&1
    echo "rc=$?"; tail -3 gpurun_out/${OUT}_plain.log; tail -2 gpurun_out/${OUT}_ncu.log ;;
  ncu-step)
    OUT=${1:-prof_step}
    python tools/quick_bench.py stick 4096 > gpurun_out/${OUT}_plain.log 2>&1 && \
    ncu --set full --clock-control none --import-source on -s 202 -c 4 -f -o gpurun_out/$OUT python tools/quick_bench.py stick 4096 > gpurun_out/${OUT}_ncu.log 2>&1
    echo "rc=$?"; tail -3 gpurun_out/${OUT}_plain.log; tail -2 gpurun_out/${OUT}_ncu.log ;;
  guards)
    python -m pytest tests/test_gpu_guards.py -q -m gpu 2>&1 | tee gpurun_out/guards.log | tail -8 ;;
  *) echo "unknown command $cmd"; exit 2 ;;
esac

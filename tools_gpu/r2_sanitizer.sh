# compute-sanitizer, ONE tool per gpurun call: usage r2_sanitizer.sh <memcheck|racecheck|synccheck|initcheck>
mkdir -p gpurun_out
TOOL=${1:-memcheck}
python tools_gpu/sanitize_case.py > gpurun_out/san_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/san_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 --log-file gpurun_out/san_${TOOL}.log python tools_gpu/sanitize_case.py > gpurun_out/san_${TOOL}_stdout.log 2>&1
echo "sanitizer rc=$?"; tail -4 gpurun_out/san_${TOOL}_stdout.log; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error" gpurun_out/san_${TOOL}.log | head -12

set -x
mkdir -p gpurun_out/san2
timeout 300 python tools/sanitize_targets.py > gpurun_out/san2/plain.log 2>&1; echo plain rc=$?; tail -3 gpurun_out/san2/plain.log
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python tools/sanitize_targets.py > gpurun_out/san2/$tool.log 2>&1; echo $tool rc=$?
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok" gpurun_out/san2/$tool.log | tail -6
  grep -E "Invalid|Race|hazard" gpurun_out/san2/$tool.log | head -10
done

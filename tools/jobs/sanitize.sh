#!/bin/bash
# compute-sanitizer logs (VERDICT r1 #9 / SURVEY 5.2): memcheck, racecheck, synccheck over tools/sanitize_probe.py
mkdir -p gpurun_out
python tools/sanitize_probe.py 2>&1 | tail -2
for tool in memcheck synccheck racecheck; do
  timeout 420 compute-sanitizer --tool $tool --print-limit 20 --log-file gpurun_out/sanitizer_$tool.log \
      python tools/sanitize_probe.py > gpurun_out/sanitizer_${tool}_stdout.log 2>&1
  echo "$tool exit $?"; tail -3 gpurun_out/sanitizer_${tool}_stdout.log; grep -c "=========" gpurun_out/sanitizer_$tool.log
  grep "ERROR SUMMARY\|RACECHECK SUMMARY" gpurun_out/sanitizer_$tool.log
done

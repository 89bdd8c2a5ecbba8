#!/bin/bash
# GPU session r02a: compute-sanitizer memcheck / racecheck / initcheck over smoke() and a 64-proof verify + prove
# (VERDICT r01 missing #8). Summaries go to gpurun_out/ and are copied to profiles/.
set -x
mkdir -p gpurun_out
SMOKE='import __graft_entry__ as g; g.smoke()'
for tool in memcheck racecheck initcheck synccheck; do
  out=gpurun_out/r02a_sanitizer_${tool}.txt
  : > $out
  for cmd in "python -c '$SMOKE'" "python tools/run_case.py verify 64 1" "python tools/run_case.py prove 64 1" "python tools/run_case.py msm 14 1"; do
    echo "==== compute-sanitizer --tool $tool : $cmd" >> $out
    eval timeout 900 compute-sanitizer --tool $tool --print-limit 20 $cmd 2>&1 | grep -v '^$' | tail -25 >> $out
    echo "rc=$?" >> $out
  done
done
grep -H "ERROR SUMMARY\|RACECHECK SUMMARY\|====" gpurun_out/r02a_sanitizer_*.txt

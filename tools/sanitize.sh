#!/bin/bash
# compute-sanitizer (memcheck, racecheck, synccheck) over the known-answer probes of every kernel family at small shapes.
# Runs on the GPU box; one log per (tool, family) under gpurun_out/sanitize/, a summary line each on stdout.
#   bash tools/sanitize.sh [families...]        default: all of tools/sanitize_target.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/sanitize
FAMS=${@:-"stream umma pair post tf32 simt lr gen"}
python tools/sanitize_target.py $FAMS > gpurun_out/sanitize/plain.log 2>&1; echo "plain rc=$? $(grep -c '^ok' gpurun_out/sanitize/plain.log) families ok"
for tool in memcheck racecheck synccheck; do
  for fam in $FAMS; do
    log=gpurun_out/sanitize/${tool}_${fam}.log
    timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python tools/sanitize_target.py $fam > $log 2>&1
    rc=$?
    echo "$tool $fam rc=$rc $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -n 1)"
  done
done

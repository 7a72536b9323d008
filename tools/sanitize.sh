#!/bin/bash
# usage (on the GPU box, via gpurun): tools/sanitize.sh <tag>  — compute-sanitizer memcheck / racecheck / synccheck / initcheck over tools/sanitize_target.py.
# Full logs stay in gpurun_out/; the summaries (every line with an error or a hazard + the ERROR SUMMARY lines) are what profiles/ keeps.
tag=$1
python tools/sanitize_target.py > gpurun_out/sanitize_plain_$tag.log 2>&1 || { echo "target fails without the sanitizer"; tail -5 gpurun_out/sanitize_plain_$tag.log; exit 1; }
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 50 python tools/sanitize_target.py > gpurun_out/sanitizer_${tool}_$tag.log 2>&1
  rc=$?
  { echo "# compute-sanitizer --tool $tool python tools/sanitize_target.py   (exit code $rc)"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|Uninitialized|Barrier error|sanitize target|smoke:" gpurun_out/sanitizer_${tool}_$tag.log | head -60
  } > gpurun_out/sanitizer_${tool}_$tag.txt
  cat gpurun_out/sanitizer_${tool}_$tag.txt
done

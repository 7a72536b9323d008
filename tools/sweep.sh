#!/bin/bash
# usage: tools/sweep.sh "<env assignments>" ...   — runs bench.py per env set and prints one summary line each
for e in "$@"; do
  env $e python bench.py --steps 24 --warmup 3 --no-cpu-baseline > gpurun_out/sweep.json 2> gpurun_out/sweep.err || { echo "$e FAILED"; tail -3 gpurun_out/sweep.err; continue; }
  python - "$e" <<'PY'
import json, sys
d=json.load(open("gpurun_out/sweep.json"))
print(sys.argv[1], "Msamples/s", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "Mrays/s", round(d["mrays_per_s"],1), {k:round(v,3) for k,v in d["stage_ms_per_step"].items()}, "nodes/ray", round(d["roofline"]["nodes_per_ray"],1))
PY
done

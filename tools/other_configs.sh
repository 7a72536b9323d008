#!/bin/bash
# usage (on the GPU box, via gpurun): tools/other_configs.sh <tag>  — the BASELINE configurations other than the headline one through bench.py
# (C4: 3840x2160 foveated sv4 preset; C3: Lost-Empire-class, 32 spp per launch; C1: Cornell box 512^2, 32 spp per launch) -> gpurun_out/<tag>_bench_c{4,3,1}.json
tag=${1:-cfg}
L="--no-cpu-baseline --no-c5 --no-tile-split --warmup 3"
python bench.py --preset sv4_fov --width 3840 --height 2160 --steps 32 $L > gpurun_out/${tag}_bench_c4.json 2>/dev/null
python bench.py --scene lost_empire --preset sv3_uniform --spl 32 --steps 8 $L > gpurun_out/${tag}_bench_c3.json 2>/dev/null
python bench.py --scene cornell --width 512 --height 512 --preset sv3_uniform --spl 32 --steps 16 $L > gpurun_out/${tag}_bench_c1.json 2>/dev/null
for c in c4 c3 c1; do python - "$tag" "$c" <<'PY'
import json, sys
tag, c = sys.argv[1], sys.argv[2]
d = json.loads(open(f"gpurun_out/{tag}_bench_{c}.json").read().strip().splitlines()[-1])
e = d["e2e"]
print(c, d["config"]["workload"][:70], "| value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "one-at-a-time", round(d["value_one_at_a_time"], 1),
      "| e2e", round(e["value"], 1), "serial", round(e["one_at_a_time_value"], 1), "moving", round(e["moving_camera_value"], 1), "async", round(e["async_value"], 1),
      "| stages", {k: round(v, 3) for k, v in d["stage_ms_per_step"].items()})
PY
done

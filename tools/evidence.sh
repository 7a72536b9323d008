#!/bin/bash
# usage (on the GPU box, via gpurun): tools/evidence.sh <tag>  — GPU tests, smoke, both bench arms, the ncu launch list, one ncu --set full capture of
# every trace / shade / gather / finalize launch of ONE frame, and one of two traceKernel launches of the 20 M-triangle 4K config.  Each profiler
# pass runs only after its own command has exited 0 without ncu.  Read the results here with tools/summarise_profile.py <tag> <tag>c5.
tag=$1
LIGHT="--steps 2 --warmup 1 --no-cpu-baseline --no-c5 --no-tile-split"
python -m pytest tests -m gpu -x -q > gpurun_out/gputests_$tag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests_$tag.log
python __graft_entry__.py smoke > gpurun_out/smoke_$tag.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || { tail -5 gpurun_out/bench_$tag.err; exit 1; }
python bench.py $LIGHT > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py $LIGHT > gpurun_out/ncu_launches_$tag.log 2>&1
# one frame = 9 traceKernel + 8 shadeKernel + finalizeKernel = 18 matched launches (the gather is fused into finalizeKernel): skip frame 1, capture frame 2
ncu --set full --clock-control none --import-source on -k regex:"traceKernel|shadeKernel|gatherKernel|finalizeKernel" -s 18 -c 18 -o gpurun_out/prof_$tag python bench.py $LIGHT > gpurun_out/ncu_full_$tag.log 2>&1
C5="--scene city --tris 20000000 --width 3840 --height 2160 $LIGHT"
python bench.py $C5 > gpurun_out/bench_c5frame_$tag.json 2> gpurun_out/bench_c5frame_$tag.err && \
ncu --set full --clock-control none -k regex:"traceKernel" -s 10 -c 2 -o gpurun_out/prof_${tag}c5 python bench.py $C5 > gpurun_out/ncu_full_${tag}c5.log 2>&1
tail -3 gpurun_out/gputests_$tag.log; tail -1 gpurun_out/smoke_$tag.log; cut -c1-300 gpurun_out/bench_$tag.json

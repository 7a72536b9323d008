#!/bin/bash
# usage (on the GPU box, via gpurun): tools/evidence.sh <tag>  — GPU tests, smoke, both bench arms, the ncu launch list and one ncu --set full capture
# of two traceKernel + two shadeKernel launches (bounces 2-3).  Each profiler pass runs only after its own command has exited 0 without ncu.
tag=$1
python -m pytest tests -m gpu -x -q > gpurun_out/gputests_$tag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests_$tag.log
python __graft_entry__.py smoke > gpurun_out/smoke_$tag.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_$tag.json 2> gpurun_out/bench_reference_$tag.err
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || exit 1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"traceKernel|shadeKernel" -s 20 -c 4 -o gpurun_out/prof_$tag python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/gputests_$tag.log; tail -1 gpurun_out/smoke_$tag.log; cut -c1-300 gpurun_out/bench_$tag.json

#!/bin/bash
# Run on the GPU box (under gpurun): plain bench run, then the ncu launch list and one full capture
# of the stage kernel.  Usage: tools/profile_stage.sh <tag>
tag=${1:-rX}
args="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-features"
python bench.py $args > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py $args > gpurun_out/ncu_launch_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:StageKernel -s 10 -c 1 -o gpurun_out/prof_$tag \
    python bench.py $args > gpurun_out/ncu_full_$tag.log 2>&1
tail -2 gpurun_out/ncu_full_$tag.log

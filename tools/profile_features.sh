#!/bin/bash
# Run on the GPU box (under gpurun): launch lists (gpu__time_duration, cold cache, serialised) of the feature
# paths next to the headline one.  Each command runs plain first, then under ncu.
for feat in "--viscous 0.01" "--limiter" "--energy-viscosity"; do
  tag=$(echo $feat | tr -d ' -.' )
  args="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-graph $feat"
  python bench.py $args > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_$tag.csv \
      python bench.py $args > gpurun_out/ncu_launch_$tag.log 2>&1
done

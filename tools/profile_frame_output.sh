#!/bin/bash
# Run on the GPU box (under gpurun): launch list (gpu__time_duration, cold cache, serialised) of one frame written by
# vtk.Writer on BASELINE config 2's mesh.  The command runs plain first, then under ncu; only the frame's kernels are
# kept in the list (the time steps before it are the headline path's launches).
args="--steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph --frame-output"
python bench.py $args > gpurun_out/plain_frame.log 2>&1 || { tail -5 gpurun_out/plain_frame.log; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'Vtk|Token|Format' -c 60 --csv --log-file gpurun_out/launches_frame.csv \
    python bench.py $args > gpurun_out/ncu_launch_frame.log 2>&1
tail -3 gpurun_out/launches_frame.csv

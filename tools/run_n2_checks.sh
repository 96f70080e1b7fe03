set -x
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $T tests/mgpu_worker.py 2 peer 2>&1 | grep MGPU > gpurun_out/r2_final_mgpu_2.log
timeout 300 $T tests/mgpu_worker.py 3 peer 2>&1 | grep MGPU >> gpurun_out/r2_final_mgpu_2.log
timeout 300 $T bench.py --gpus 2 --steps 300 --warmup 3 > gpurun_out/r2_final_n2.json 2> gpurun_out/r2_final_n2.err; echo rc $?
timeout 300 $T bench.py --gpus 2 --steps 60 --warmup 3 --degree 3 --cells 32 > gpurun_out/r2_final_n2_p3.json 2> gpurun_out/r2_final_n2_p3.err; echo rc $?
timeout 300 $T tools/trace_step.py > gpurun_out/r2_final_trace_n2.txt 2> gpurun_out/r2_final_trace_n2.err; echo rc $?
cat gpurun_out/r2_final_mgpu_2.log; tail -c 300 gpurun_out/r2_final_n2.err; tail -c 300 gpurun_out/r2_final_n2_p3.err

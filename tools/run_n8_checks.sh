# 8 GPUs of one box: parity of the device-side exchange against the single-part oracle, then the weak-scaling bench line
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $T tests/mgpu_worker.py 2 peer 2>&1 | grep MGPU > gpurun_out/r2_final_mgpu_8.log
timeout 300 $T bench.py --gpus 8 --steps 300 --warmup 3 > gpurun_out/r2_final_n8.json 2> gpurun_out/r2_final_n8.err; echo rc $?
cat gpurun_out/r2_final_mgpu_8.log; tail -c 300 gpurun_out/r2_final_n8.err

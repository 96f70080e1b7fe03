# A/B of the two p = 2 stage kernels on the GPU box: parity tests and the bench with MCFD_STAGE_KERNEL=nodes, then lines
export MCFD_STAGE_KERNEL=nodes
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py tests/test_gpu_peer.py -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 60 --warmup 3 --no-cpu-baseline --no-e2e --no-features > gpurun_out/nodes_n1.json 2> gpurun_out/nodes_n1.err; echo rc $?; tail -c 400 gpurun_out/nodes_n1.err
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/nodes_n1.json").read().strip().splitlines()[-1])
    print("nodes", d["ms_per_step"], d["value"], d["roofline"]["kernel_ms_per_stage"], d["roofline"]["frac"], d["check"])
except Exception as e: print("no line", e)
PY

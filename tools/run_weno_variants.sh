# A/B of the resident-CTA target of the limiter kernels: build_variants/lib_weno_<minb>.so copied over the library in turn
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-graph --limiter"   # add "--degree 3 --cells 32" for the p = 3 kernels
cp minicfd_b200/libminicfd_b200.so /tmp/orig.so
for f in build_variants/lib_weno_*.so; do
  v=$(basename $f .so)
  cp $f minicfd_b200/libminicfd_b200.so
  $B > gpurun_out/$v.json 2> gpurun_out/$v.err || { echo "$v failed"; tail -2 gpurun_out/$v.err; continue; }
  python - <<PY
import json
d=json.loads(open("gpurun_out/$v.json").read().strip().splitlines()[-1])
print("$v", round(d["ms_per_step"],4), "ms/step; stage kernel", round(d["roofline"]["kernel_ms_per_stage"],4))
PY
done
cp /tmp/orig.so minicfd_b200/libminicfd_b200.so
python -m pytest tests/test_gpu_limiter.py -q 2>&1 | tail -2

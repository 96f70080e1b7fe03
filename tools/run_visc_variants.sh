# A/B of launch shapes of the viscous kernels: build_variants/lib_*.so are copied over the library in turn.
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-graph --viscous 0.01"
cp minicfd_b200/libminicfd_b200.so /tmp/orig.so
for f in build_variants/lib_v_*.so; do
  v=$(basename $f .so)
  cp $f minicfd_b200/libminicfd_b200.so
  $B > gpurun_out/visc_$v.json 2> gpurun_out/visc_$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/visc_$v.json").read().strip().splitlines()[-1])
s=d["roofline"]["kernel_ms_per_stage"]; t=d["ms_per_step"]
print("$v", round(t,4), "ms/step; stage kernel", round(s,4), "trace ~", round((t-3*s)/3,4))
PY
done
cp /tmp/orig.so minicfd_b200/libminicfd_b200.so

# A/B of launch shapes of the p = 3 stage kernel: build_variants/lib_p3_<cells>_<warps>_<minb>.so copied over the library
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --degree 3 --cells 32"
cp minicfd_b200/libminicfd_b200.so /tmp/orig.so
for f in build_variants/lib_p3_*.so; do
  v=$(basename $f .so)
  cp $f minicfd_b200/libminicfd_b200.so
  $B > gpurun_out/$v.json 2> gpurun_out/$v.err || { echo "$v failed"; tail -2 gpurun_out/$v.err; continue; }
  python - <<PY
import json
d=json.loads(open("gpurun_out/$v.json").read().strip().splitlines()[-1])
print("$v", round(d["ms_per_step"],4), "ms/step; stage kernel", round(d["roofline"]["kernel_ms_per_stage"],4), round(d["value"]/1e9,2), "G", round(d["roofline"]["frac"],4))
PY
done
cp /tmp/orig.so minicfd_b200/libminicfd_b200.so

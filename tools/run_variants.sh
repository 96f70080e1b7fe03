#!/bin/bash
# On the GPU box: time bench.py with every build_variants/lib_*.so in place of the library.
# Usage: tools/run_variants.sh [extra bench.py args]
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e $*"
cp minicfd_b200/libminicfd_b200.so /tmp/orig.so
for f in build_variants/lib_*.so; do
  v=$(basename $f .so)
  cp $f minicfd_b200/libminicfd_b200.so
  $B > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err || { echo "$v failed"; tail -2 gpurun_out/var_$v.err; continue; }
  python - <<PY
import json
d=json.loads(open("gpurun_out/var_$v.json").read().strip().splitlines()[-1])
print("$v", round(d["ms_per_step"],4), "ms/step; stage kernel", round(d["roofline"]["kernel_ms_per_stage"],4), "frac", round(d["roofline"]["frac"],3))
PY
done
cp /tmp/orig.so minicfd_b200/libminicfd_b200.so

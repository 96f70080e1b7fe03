# A/B of launch shapes of the headline stage kernel: build_variants/lib_<cells>_<warps>_<minb>.so are copied over the
# library in turn (each is the whole library compiled with -DMCFD_P2_CELLS/-DMCFD_P2_WARPS/-DMCFD_P2_MINB).
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
cp minicfd_b200/libminicfd_b200.so /tmp/orig.so
$B > gpurun_out/stage_orig.json 2> gpurun_out/stage_orig.err
python -c "
import json
d=json.loads(open('gpurun_out/stage_orig.json').read().strip().splitlines()[-1])
print('orig', round(d['ms_per_step'],4), 'ms/step; stage kernel', round(d['roofline']['kernel_ms_per_stage'],4))"
for f in build_variants/lib_*.so; do
  v=$(basename $f .so)
  cp $f minicfd_b200/libminicfd_b200.so
  $B > gpurun_out/stage_$v.json 2> gpurun_out/stage_$v.err || { echo "$v failed"; tail -2 gpurun_out/stage_$v.err; continue; }
  python - <<PY
import json
d=json.loads(open("gpurun_out/stage_$v.json").read().strip().splitlines()[-1])
print("$v", round(d["ms_per_step"],4), "ms/step; stage kernel", round(d["roofline"]["kernel_ms_per_stage"],4))
PY
done
cp /tmp/orig.so minicfd_b200/libminicfd_b200.so

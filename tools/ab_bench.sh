mkdir -p gpurun_out
for rep in 1 2; do
for lib in default build_ab/libsrt_item1.so; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  python bench.py --no-cpu-baseline > gpurun_out/abb.log 2>&1
  python - <<'P'
import json
for l in open('gpurun_out/abb.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['value_one_piece']['value'], d['stage_ms_one_piece']['shadow'], d['e2e']['value'], d['c3_frame0']['value'])
P
done; done

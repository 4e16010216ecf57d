#!/bin/bash
# Rebuilds libtinyr.so with different tile shapes / CTA sizes (of the lone-frame shape; the batch shape has its own
# -DTR_RASTER_THREADS_BATCH / -DTR_LIST_CAP_BATCH / -DTR_MIN_CTAS_*_BATCH knobs, see tools/tune_ao.sh) on the GPU box and
# prints the bench numbers.
# usage (under gpurun): bash tools/tune_raster.sh "256 64 32" "512 64 32" ...
set -u
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  set -- $cfg
  rt=$1; tw=$2; th=$3; xf="${4:-}"
  touch tiny_renderer_b200/csrc/tinyr.cu
  make -s -C tiny_renderer_b200/csrc EXTRA="-DTR_RASTER_THREADS=$rt -DTR_TILE_W=$tw -DTR_TILE_H=$th $xf" all > /dev/null 2>&1 || { echo "build failed $cfg"; continue; }
  python bench.py --steps 40 --warmup 3 > gpurun_out/tune_${rt}_${tw}_${th}.log 2>&1
  python - "$rt $tw $th" gpurun_out/tune_${rt}_${tw}_${th}.log <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    k=d["kernel_us"]
    print("cfg", sys.argv[1], "value %.0f" % d["value"], "phong K1 %.1f K3 %.1f | darboux K1 %.1f K3 %.1f" % (k["phong"]["K1_vertex_setup"],k["phong"]["K3_raster_visibility"],k["darboux"]["K1_vertex_setup"],k["darboux"]["K3_raster_visibility"]), "ao %.0f" % d["ao"]["value"])
except Exception as e:
    print("cfg", sys.argv[1], "failed", e); print(open(sys.argv[2]).read()[-1500:])
PY
done
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

#!/bin/bash
# Rebuilds libtinyr.so with different -D flags on the GPU box and prints the AO kernel times and the AO bench value.
# usage (under gpurun): bash tools/tune_ao.sh "-DTR_AO_GROUP=4" "-DTR_AO_GROUP=8 -DTR_AO_MIN_CTAS=8" ...
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for xf in "$@"; do
  i=$((i+1))
  touch tiny_renderer_b200/csrc/tinyr.cu
  make -s -C tiny_renderer_b200/csrc EXTRA="$xf" all > /dev/null 2>&1 || { echo "build failed: $xf"; continue; }
  grep -A2 "k_ao_accumulate" tiny_renderer_b200/csrc/ptxas_tinyr.log | grep -o "Used [0-9]* registers"
  echo "cfg [$xf] $(python tools/ao_profile.py 2>&1 | tail -1)"
  python bench.py --steps 20 --warmup 3 > gpurun_out/tune_ao_$i.log 2>&1
  python - gpurun_out/tune_ao_$i.log <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("   value %.0f  ao %.0f passes/s checksum %d" % (d["value"], d["ao"]["value"], d["ao"]["checksum"]))
except Exception as e:
    print("   failed", e); print(open(sys.argv[1]).read()[-800:])
PY
done
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

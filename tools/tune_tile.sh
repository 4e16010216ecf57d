#!/bin/bash
# Rebuilds libtinyr.so with different -D flags on the GPU box; per build: lone-frame device times (tools/lone_frame_time.py)
# and the 4K view-batch throughput (tools/batch_probe.py).  Restores the product build at the end.
# usage (under gpurun): bash tools/tune_tile.sh "" "-DTR_TILE_H=16" ...
set -u
cd "$(dirname "$0")/.."
for xf in "$@"; do
  touch tiny_renderer_b200/csrc/tinyr.cu
  make -s -C tiny_renderer_b200/csrc EXTRA="$xf" all > /dev/null 2>&1 || { echo "build failed: $xf"; continue; }
  echo "cfg [$xf] $(python tools/lone_frame_time.py 2>&1 | tail -1)"
  echo "cfg [$xf] $(PROBE_REPS=${PROBE_REPS:-3} python tools/batch_probe.py 2>&1 | tail -1)"
done
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

#!/bin/bash
# Rebuilds libtinyr.so with different -D flags on the GPU box and prints the lone-frame device times of the single-frame
# BASELINE configs (tools/lone_frame_time.py: clear / set_* / render through the C ABI, CUDA events, median of 31).
# usage (under gpurun): bash tools/tune_lone.sh "" "-DTR_RASTER_THREADS=512 -DTR_MIN_CTAS_LIGHT=3 -DTR_MIN_CTAS_DARBOUX=2" ...
set -u
cd "$(dirname "$0")/.."
for xf in "$@"; do
  touch tiny_renderer_b200/csrc/tinyr.cu
  make -s -C tiny_renderer_b200/csrc EXTRA="$xf" all > /dev/null 2>&1 || { echo "build failed: $xf"; continue; }
  echo "cfg [$xf] $(python tools/lone_frame_time.py 2>&1 | tail -1)"
done
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

#!/bin/bash
# Builds the TR_TIMELINE diagnostic variant of libtinyr.so on the GPU box, runs tools/lone_timeline.py, restores the product build.
# usage (under gpurun): bash tools/lone_timeline.sh [pipeline ...]  > gpurun_out/timeline.jsonl     (TR_EXTRA adds -D flags)
set -u
cd "$(dirname "$0")/.."
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc EXTRA="-DTR_TIMELINE ${TR_EXTRA:-}" all > /dev/null 2>&1 || { echo "build failed"; exit 1; }
python tools/lone_timeline.py "$@"
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

#!/bin/bash
# usage (under gpurun): bash tools/lone_order_probe.sh [pipeline ...] > gpurun_out/order_probe.jsonl
set -u
cd "$(dirname "$0")/.."
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc EXTRA="-DTR_TIMELINE" all > /dev/null 2>&1 || { echo "build failed"; exit 1; }
python tools/lone_order_probe.py "$@"
touch tiny_renderer_b200/csrc/tinyr.cu
make -s -C tiny_renderer_b200/csrc all > /dev/null 2>&1

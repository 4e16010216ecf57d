#!/bin/bash
# A/B of the lone-frame launch path on one box: the committed baseline build (_ab/base, git worktree of HEAD built there)
# against the working tree with its switches.  usage (under gpurun): bash tools/ab_lone.sh > gpurun_out/ab_lone.txt
set -u
cd "$(dirname "$0")/.."
B=$PWD/_ab/base/tiny_renderer_b200/libtinyr.so
for i in 1 2; do
[ -f "$B" ] && echo "base                $(TINYR_LIB=$B python tools/lone_frame_time.py 2>&1 | tail -1)"
echo "sorted order        $(python tools/lone_frame_time.py 2>&1 | tail -1)"
echo "TINYR_LONE_SORT=0   $(TINYR_LONE_SORT=0 python tools/lone_frame_time.py 2>&1 | tail -1)"
done

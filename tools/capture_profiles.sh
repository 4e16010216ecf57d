#!/bin/bash
# Captures the evidence kept under profiles/ for one build, on the GPU box (run under gpurun):
#   1. the bench line and the reference arm, no profiler attached;
#   2. the launch list of the same bench command (gpu__time_duration per launch);
#   3. `ncu --set full` of the tile kernels (batch shape, then the lone-frame shape) / vertex set-up from the bench and
#      of the AO kernels from tools/ao_profile.py.
# Everything lands in gpurun_out/ as <tag>_*; tools/ncu_summary.py turns the .ncu-rep files into the CSV / JSON
# summaries that are committed.  usage: bash tools/capture_profiles.sh r1e
set -u
cd "$(dirname "$0")/.."
tag="${1:-cap}"
out=gpurun_out
mkdir -p $out
python bench.py --steps 100 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -5 $out/${tag}_bench.err; exit 1; }
tail -c 400 $out/${tag}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err || echo "reference arm failed"
python bench.py --steps 2 --warmup 3 > $out/${tag}_plain.log 2>&1 || { echo "short bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_raster|k_vertex_setup" -s 8 -c 12 -f -o $out/${tag}_full \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu2.log 2>&1
# the 256-thread shape of the tile kernel only runs in the bench's one-frame-at-a-time section (24 phong frames, then darboux)
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_raster<.*256>" -s 22 -c 4 -f -o $out/${tag}_full_lone \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu4.log 2>&1
python tools/ao_profile.py > $out/${tag}_ao_plain.log 2>&1 || { echo "ao_profile failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"k_ao|k_raster" -s 20 -c 4 -f -o $out/${tag}_full_ao \
    env AO_PROFILE_RING=2 python tools/ao_profile.py > $out/${tag}_ncu3.log 2>&1
ls -la $out/${tag}_*

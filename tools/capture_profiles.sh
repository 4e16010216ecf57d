#!/bin/bash
# Captures the evidence kept under profiles/ for one build, on the GPU box (run under gpurun):
#   1. the bench line and the reference arm, no profiler attached;
#   2. the launch list of a short bench run (gpu__time_duration per launch; TINYR_BENCH_QUICK=1: one turn of the
#      64-angle turntable per step, 32 AO lights, 64 8K views — same kernels, fewer launches);
#   3. dram__bytes per launch over 256 consecutive tile-kernel launches of the de-aliased view batch with L2 contents
#      carried over (--cache-control none): the steady-state DRAM traffic per frame;
#   4. `ncu --set full` of the tile kernels (batch shape, then the lone-frame shape) / vertex set-up / host delivery from
#      the bench and of the AO kernels from tools/ao_profile.py.
# Everything lands in gpurun_out/ as <tag>_*; tools/ncu_summary.py turns the .ncu-rep files into the CSV / JSON
# summaries that are committed.  usage: bash tools/capture_profiles.sh r2a [skipbench|loneonly|part1|part2]
# (gpurun brings back at most 64 MiB per call: part1 = bench, launch list, steady traffic, batch + host captures; part2 = lone + AO captures)
set -u
cd "$(dirname "$0")/.."
tag="${1:-cap}"
out=gpurun_out
mkdir -p $out
mode="${2:-}"
if [ "$mode" != "skipbench" ] && [ "$mode" != "loneonly" ] && [ "$mode" != "part2" ]; then
  python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -5 $out/${tag}_bench.err; exit 1; }
  tail -c 300 $out/${tag}_bench.json
  python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err || echo "reference arm failed"
fi
export TINYR_BENCH_QUICK=1
lone_captures() {
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_raster<.*Phong, .*256>|k_vertex_setup<.*0, .*1>" -s 40 -c 8 -f -o $out/${tag}_full_lone \
      python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu4.log 2>&1
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_raster<.*Darboux, .*256>" -s 20 -c 6 -f -o $out/${tag}_full_lone2 \
      python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu7.log 2>&1
}
ao_captures() {
  python tools/ao_profile.py > $out/${tag}_ao_plain.log 2>&1 || { echo "ao_profile failed"; exit 1; }
  ncu --set full --clock-control none --import-source on -k regex:"k_ao|k_raster|k_vertex_setup" -s 20 -c 6 -f -o $out/${tag}_full_ao \
      env AO_PROFILE_RING=2 python tools/ao_profile.py > $out/${tag}_ncu3.log 2>&1
}
if [ "$mode" = "loneonly" ]; then lone_captures; ls -la $out/${tag}_full_lone*; exit 0; fi
if [ "$mode" = "part2" ]; then lone_captures; ao_captures; ls -la $out/${tag}_*; exit 0; fi
python bench.py --steps 2 --warmup 3 > $out/${tag}_plain.log 2>&1 || { echo "short bench failed"; tail -5 $out/${tag}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 900 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu1.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none --clock-control none \
    --kernel-name-base demangled -k regex:"k_raster" -s 420 -c 256 --csv --log-file $out/${tag}_steady.csv \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu5.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_raster<.*128>|k_vertex_setup" -s 120 -c 16 -f -o $out/${tag}_full \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu2.log 2>&1
# the 256-thread shape of the tile kernel only runs in the bench's one-frame-at-a-time sections; one capture per pipeline
# (the sections of the two pipelines follow each other)
if [ "$mode" != "part1" ]; then lone_captures; fi
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_frame_to_host" -s 300 -c 3 -f -o $out/${tag}_full_host \
    python bench.py --steps 2 --warmup 3 > $out/${tag}_ncu6.log 2>&1
if [ "$mode" != "part1" ]; then ao_captures; fi
ls -la $out/${tag}_*

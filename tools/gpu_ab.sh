#!/bin/bash
# A/B of kernel variants (path-tracer_b200/ab/libpt_*.so) with tools/perf_probe.py; results -> gpurun_out/$OUT/ab.jsonl
OUT=${OUT:-ab}
mkdir -p gpurun_out/$OUT
: > gpurun_out/$OUT/ab.jsonl
for wl in "console_1080p_128spp_5b 16" "city_4k_1024spp_8b 4" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for v in $VARIANTS; do
    echo "{\"workload\": \"$1\", \"lib\": \"$v\"}" >> gpurun_out/$OUT/ab.jsonl
    PT_B200_VARIANT=$v PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_$v.so timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/$OUT/ab.jsonl 2>> gpurun_out/$OUT/ab.err
  done
done

mkdir -p gpurun_out/s12
: > gpurun_out/s12/leaf.jsonl
for wl in "console_1080p_128spp_5b 16" "city_4k_1024spp_8b 4" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for lt in 4 2 3 6 8; do
    echo "{\"workload\": \"$1\", \"lib\": \"leaf$lt\"}" >> gpurun_out/s12/leaf.jsonl
    PT_B200_LEAF_TRIS=$lt PT_B200_VARIANT=leaf$lt timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/s12/leaf.jsonl 2>> gpurun_out/s12/leaf.err
  done
done

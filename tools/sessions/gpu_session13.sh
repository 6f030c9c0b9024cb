mkdir -p gpurun_out/s13
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s13/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s13/pytest.log
: > gpurun_out/s13/leaf.jsonl
for wl in "console_1080p_128spp_5b 16" "city_4k_1024spp_8b 4" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for cfg in "sah 2" "sah 1" "lbvh 2" "lbvh 1" "lbvh 4"; do
    set -- $1 $2 $cfg
    echo "{\"workload\": \"$1\", \"lib\": \"$3-leaf$4\"}" >> gpurun_out/s13/leaf.jsonl
    PT_B200_ACCEL=$3 PT_B200_LEAF_TRIS=$4 timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/s13/leaf.jsonl 2>> gpurun_out/s13/leaf.err
  done
done

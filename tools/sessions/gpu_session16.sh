mkdir -p gpurun_out/s16
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s16/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s16/pytest.log
: > gpurun_out/s16/probe.jsonl
for wl in "console_1080p_128spp_5b 16" "city_4k_1024spp_8b 4" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for cfg in "sah" "lbvh"; do
    echo "{\"workload\": \"$1\", \"lib\": \"$cfg\"}" >> gpurun_out/s16/probe.jsonl
    PT_B200_ACCEL=$cfg timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/s16/probe.jsonl 2>> gpurun_out/s16/probe.err
  done
done
TREES=sah:2,sah:8,lbvh:2 timeout 900 python tools/tree_diff.py city_4k_1024spp_8b 3000000 > gpurun_out/s16/city_diff.jsonl 2> gpurun_out/s16/city_diff.err

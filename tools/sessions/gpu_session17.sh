mkdir -p gpurun_out/s17
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s17/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s17/pytest.log
: > gpurun_out/s17/probe.jsonl
for wl in "console_1080p_128spp_5b 16" "city_4k_1024spp_8b 4" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for cfg in "sah 0" "sah 1" "sah 0" "sah 1" "lbvh 0" "lbvh 1"; do
    set -- $1 $2 $cfg
    echo "{\"workload\": \"$1\", \"lib\": \"$3-margin$4\"}" >> gpurun_out/s17/probe.jsonl
    PT_B200_ACCEL=$3 PT_B200_MARGIN_SCALE=$4 timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/s17/probe.jsonl 2>> gpurun_out/s17/probe.err
  done
done

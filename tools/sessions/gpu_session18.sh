mkdir -p gpurun_out/s18
timeout 1500 python -m pytest tests -m gpu -x -q -k "edge_grazing or lbvh or isfb or brute" > gpurun_out/s18/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s18/pytest.log
: > gpurun_out/s18/probe.jsonl
for wl in "console_1080p_128spp_5b 16"; do
  set -- $wl
  for cfg in "sah 0" "sah 1" "sah 1.5" "sah 0" "sah 1" "sah 1.5" "lbvh 0" "lbvh 1"; do
    set -- $1 $2 $cfg
    echo "{\"workload\": \"$1\", \"lib\": \"$3-margin$4\"}" >> gpurun_out/s18/probe.jsonl
    PT_B200_ACCEL=$3 PT_B200_MARGIN_SCALE=$4 timeout 300 python tools/perf_probe.py $1 $2 >> gpurun_out/s18/probe.jsonl 2>> gpurun_out/s18/probe.err
  done
done
TREES=sah:2,sah:8,lbvh:2 timeout 900 python tools/tree_diff.py city_4k_1024spp_8b 3000000 > gpurun_out/s18/city_diff.jsonl 2> gpurun_out/s18/city_diff.err

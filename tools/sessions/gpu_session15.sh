mkdir -p gpurun_out/s15
TREES=sah:2,sah:8,sah:4 timeout 900 python tools/tree_diff.py city_4k_1024spp_8b 3000000 > gpurun_out/s15/city.jsonl 2> gpurun_out/s15/city.err

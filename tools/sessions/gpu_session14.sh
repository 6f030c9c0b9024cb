mkdir -p gpurun_out/s14
timeout 900 python tools/tree_diff.py city_4k_1024spp_8b 1048576 > gpurun_out/s14/city.jsonl 2> gpurun_out/s14/city.err
TREES=sah:2,lbvh:2,sah:8 timeout 600 python tools/tree_diff.py foliage_1080p_256spp_4b 1048576 > gpurun_out/s14/foliage.jsonl 2> gpurun_out/s14/foliage.err
TREES=sah:2,lbvh:2,sah:8 timeout 600 python tools/tree_diff.py console_1080p_128spp_5b 1048576 > gpurun_out/s14/console.jsonl 2> gpurun_out/s14/console.err

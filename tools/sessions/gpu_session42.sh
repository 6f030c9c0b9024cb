# 4 GPUs in ONE process: the cross-device leg of pt_render_multi (peer copy + ordered add) and its wall-clock scaling
R=gpurun_out/r06
mkdir -p $R
nvidia-smi topo -m > $R/topo_4gpu.txt 2>&1
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_multi or sample_partition or cli_render" > $R/pytest_4gpu.log 2>&1; echo "rc=$?" >> $R/pytest_4gpu.log
timeout 200 python tools/multi_inproc.py console_1080p_128spp_5b 128 3 > $R/multi_inproc_4gpu.json 2> $R/multi_inproc.err
tail -n 3 $R/pytest_4gpu.log; cat $R/multi_inproc_4gpu.json; tail -n 3 $R/multi_inproc.err

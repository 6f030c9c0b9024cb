R=gpurun_out/r06
mkdir -p $R
( time timeout 500 python -m pytest tests -m gpu -x -q --durations=8 ) > $R/pytest.log 2>&1; echo "pytest rc=$?" >> $R/pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $R/smoke.log 2>&1
timeout 300 python bench.py > $R/bench_console_1080p_128spp.json 2> $R/bench_console.err
# in-process multi-replica render through the CLI (one device here: --gpus all == 1), with its stats file
printf 'resolution:\n  width: 1920\n  height: 1080\nsamples: 32\nbounces: 5\n' > /tmp/p.yml
( time ./path-tracer_b200/path-tracer render tests/golden/scenes/head/scene.isf -o $R/head_1080p_32spp.png -p /tmp/p.yml --gpus all --stats $R/head_cli_stats.json ) > $R/cli_head.log 2>&1
tail -3 $R/pytest.log; cat $R/smoke.log | tail -2; cat $R/bench_console_1080p_128spp.json | cut -c1-400

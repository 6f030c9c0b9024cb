R=gpurun_out/r05
mkdir -p $R
timeout 300 python bench.py > $R/bench_console_1080p_128spp.json 2> $R/bench_console.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $R/bench_reference.json 2> $R/bench_reference.err
( time timeout 500 python -m pytest tests -m gpu -x -q --durations=12 ) > $R/pytest.log 2>&1; echo "pytest rc=$?" >> $R/pytest.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_launches.log 2>&1
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $R/smoke.log 2>&1

# First GPU session of the next round: what round 1 could not re-measure after the ordered hit list became the default.
#   build the A/B pair first:  bash tools/build_variant.sh base -DPTB_HIT_LIST=0; bash tools/build_variant.sh hitlist -DPTB_HIT_LIST=1
#   run:                        gpurun --timeout 900 -- 'bash tools/sessions/r2_first.sh'
R=gpurun_out/r07
mkdir -p $R
( time timeout 300 python -m pytest tests -m gpu -x -q --durations=8 ) > $R/pytest.log 2>&1; echo "pytest rc=$?" >> $R/pytest.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $R/smoke.log 2>&1
timeout 300 python bench.py > $R/bench_console_1080p_128spp.json 2> $R/bench_console.err
timeout 300 python bench.py --workload foliage_1080p_256spp_4b --steps 2 --warmup 3 --cpu-seconds 8 > $R/bench_foliage_256spp.json 2> $R/bench_foliage.err
timeout 400 python bench.py --workload city_4k_1024spp_8b --steps 2 --warmup 3 --no-cpu-baseline > $R/bench_city_1024spp.json 2> $R/bench_city.err
# city with and without the hit list (the one workload the round-1 A/B did not cover)
if [ -f path-tracer_b200/ab/libpt_base.so ]; then
  for v in base hitlist; do
    PT_B200_VARIANT=$v PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_$v.so timeout 200 python tools/perf_probe.py city_4k_1024spp_8b 4 >> $R/ab_hitlist_city.jsonl 2>> $R/ab.err
  done
fi
# launch list and one full capture of the kernels the hit list changed (only after the plain runs above exited)
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $R/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_shade|k_hit_list' -s 12 -c 12 -o $R/console_shade_full -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_full.log 2>&1
tail -n 3 $R/pytest.log; cut -c1-160 $R/bench_console_1080p_128spp.json

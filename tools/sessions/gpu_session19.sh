R=gpurun_out/r03
mkdir -p $R
timeout 1500 python -m pytest tests -m gpu -x -q > $R/pytest.log 2>&1; echo "pytest rc=$?" >> $R/pytest.log
timeout 900 python bench.py > $R/bench_console_1080p_128spp.json 2> $R/bench_console.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $R/bench_reference_arm.json 2> $R/bench_reference.err
timeout 900 python bench.py --workload foliage_1080p_256spp_4b --spp 64 --steps 2 --warmup 3 --cpu-seconds 8 > $R/bench_foliage_64spp.json 2> $R/bench_foliage.err
timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > $R/bench_city_16spp.json 2> $R/bench_city.err
PT_B200_ACCEL=lbvh timeout 1200 python bench.py --workload city_4k_1024spp_8b --spp 16 --steps 2 --warmup 3 --no-cpu-baseline > $R/bench_city_16spp_lbvh.json 2> $R/bench_city_lbvh.err
timeout 900 python tools/isfb_probe.py city_4k_1024spp_8b /tmp > $R/isfb_city.json 2> $R/isfb_city.err
timeout 300 python - > $R/lbvh_build_twice.json 2> $R/lbvh_build_twice.err <<'PY'
import json, sys
sys.path.insert(0, ".")
import __graft_entry__
pkg = __graft_entry__.load_package()
pkg.set_accel_builder("lbvh")
out = []
for i in range(3):
    sc = pkg.Scene.synthetic("city", 20000000, 5)
    out.append(sc.accel_info()["build_ms"])
    sc.close()
print(json.dumps(dict(lbvh_build_ms_city_20M=out)))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/launches_console_8spp.csv python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -s 18 -c 18 -o $R/console_full -f python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline > $R/ncu_full.log 2>&1

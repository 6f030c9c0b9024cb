R=gpurun_out/s37
mkdir -p $R
for v in shade6 shade7 shade8; do
  for wl in "foliage_1080p_256spp_4b 64" "console_1080p_128spp_5b 128" "city_4k_1024spp_8b 16"; do
    set -- $wl
    PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_$v.so timeout 600 python bench.py --workload $1 --spp $2 --steps 2 --warmup 2 --no-cpu-baseline > $R/${v}_$1.json 2>> $R/err.log
    python - $R/${v}_$1.json $v $1 <<'PY' >> $R/summary.txt
import json, sys
d = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")][0]
k = d["roofline"]["kernel_ms_per_step"]
print(sys.argv[2], sys.argv[3][:8], "value", round(d["value"], 1), {a: round(b, 2) for a, b in k.items()})
PY
  done
done

R=gpurun_out/s31
mkdir -p $R
for wp in 67108864 134217728 268435456 67108864 134217728 268435456; do
  PT_B200_WAVE_PATHS=$wp timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > $R/console_$wp.json 2>> $R/err.log
  python - $R/console_$wp.json $wp <<'PY' >> $R/summary.txt
import json, sys
d = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")][0]
k = d["roofline"]["kernel_ms_per_step"]
print("console wave", sys.argv[2], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), {a: round(b, 2) for a, b in k.items()}, "launches", d["gpu_launches"])
PY
done

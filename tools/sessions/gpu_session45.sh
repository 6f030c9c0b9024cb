# A/B of the ordered hit list (PTB_HIT_LIST, DESIGN section 8 first open item): console 16 spp, foliage 8 spp
OUT=gpurun_out/r06
mkdir -p $OUT
: > $OUT/ab_hitlist.jsonl
for wl in "console_1080p_128spp_5b 16" "foliage_1080p_256spp_4b 8"; do
  set -- $wl
  for v in base hitlist; do
    echo "{\"workload\": \"$1\", \"lib\": \"$v\"}" >> $OUT/ab_hitlist.jsonl
    PT_B200_VARIANT=$v PT_B200_LIB=$PWD/path-tracer_b200/ab/libpt_$v.so timeout 40 python tools/perf_probe.py $1 $2 >> $OUT/ab_hitlist.jsonl 2>> $OUT/ab_hitlist.err
  done
done
cat $OUT/ab_hitlist.jsonl; tail -n 3 $OUT/ab_hitlist.err

# bench line of the round's final library (launch-geometry caches are on the launch path of every kernel)
R=gpurun_out/r06
mkdir -p $R
timeout 80 python bench.py --no-cpu-baseline > $R/bench_console_final.json 2> $R/bench_console_final.err
cut -c1-200 $R/bench_console_final.json

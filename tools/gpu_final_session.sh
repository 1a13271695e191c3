TAG=r01o
OUT=/tmp/prof_$TAG; mkdir -p $OUT gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "ref rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv python bench.py --vrls 2000 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "launch list rc=$?"
cp $OUT/* gpurun_out/ 2>/dev/null
tail -c 2500 $OUT/bench_$TAG.json

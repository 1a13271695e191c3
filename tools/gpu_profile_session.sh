#!/bin/bash
# One GPU session: full bench (plain), the reference arm, then the launch list and an ncu --set full capture of the
# no-occluder gather kernel.  Long-running processes write to /tmp; results are copied into gpurun_out/ at the end.
TAG=${1:-r01}
OUT=/tmp/prof_$TAG; mkdir -p $OUT gpurun_out
timeout 600 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
tail -c 3000 $OUT/bench_$TAG.json
timeout 300 python bench.py --impl reference > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "reference arm rc=$?"
tail -c 1200 $OUT/bench_ref_$TAG.json
SMALL="--vrls 2000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-secondary"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv python bench.py --vrls 2000 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gather_fast_kernel -c 1 -f -o $OUT/${TAG}_gather_v1 python bench.py $SMALL --no-occluders > $OUT/ncu_v1.log 2>&1; echo "ncu v1 rc=$?"
cp $OUT/* gpurun_out/ 2>/dev/null
ls -la gpurun_out | tail -8

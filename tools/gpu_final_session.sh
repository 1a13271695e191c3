#!/bin/bash
# Round-end measurement set on one B200: smoke, bench (both arms), the ncu launch list of the bench command.
TAG=${1:-r02i}
OUT=/tmp/prof_$TAG; mkdir -p $OUT gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref rc=$?"
# launch list of the SAME command at a reduced VRL count (ncu serialises and replays; the kernel's share of the step is what must agree)
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv python bench.py --vrls 8000 --steps 2 --warmup 1 --no-cpu-baseline --no-secondary > $OUT/ncu_launch.log 2>&1; echo "launch list rc=$?"
cp $OUT/* gpurun_out/ 2>/dev/null
tail -c 3000 $OUT/${TAG}_bench_n1.json

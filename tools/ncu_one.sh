#!/bin/bash
# ncu --set full capture of one gather kernel: tools/ncu_one.sh <tag> <kernel regex> [extra bench flags]
TAG=$1; KRE=$2; shift 2
OUT=/tmp/ncu_$TAG; mkdir -p $OUT gpurun_out
SMALL="--vrls 2000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-secondary"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:$KRE -c 1 -f -o $OUT/$TAG python bench.py $SMALL "$@" > $OUT/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
cp $OUT/* gpurun_out/; ls -la gpurun_out/$TAG.ncu-rep

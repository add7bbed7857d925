#!/bin/bash
mkdir -p gpurun_out
T=gpurun_out/ab_$1.txt
run() { local b=$1 na=$2; shift 2; env "$@" python scripts/dev_bwd_ext_time.py $b $na 2>&1 | tail -1; }
{
for shape in "128 1024" "256 1024" "512 1024"; do
  set -- $shape
  echo "== $1 rows x $2 audio, sharded-rank launch (dv on its own stream)"
  run $1 $2 AVT_X=rule
  run $1 $2 AVT_DV_PAIRS=40
  run $1 $2 AVT_DV_PAIRS=36
  run $1 $2 AVT_DV_PAIRS=30
  run $1 $2 AVT_DV_PAIRS=24
  run $1 $2 AVT_NO_COSCHED=1
done
} > $T 2>&1
cat $T

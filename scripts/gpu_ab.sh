#!/bin/bash
# dev: check of the co-scheduling rule at small shapes (one process per env-switch variant)
mkdir -p gpurun_out
T=gpurun_out/ab_$1.txt
run() { local b=$1 na=$2; shift 2; env "$@" python scripts/dev_bwd_time.py $b $na 2>&1 | tail -1; }
{
for shape in "128 128" "256 256" "512 512" "1024 1024" "128 1024" "64 1024"; do
  set -- $shape
  echo "== $1 rows x $2 audio"
  run $1 $2 AVT_X=rule
  run $1 $2 AVT_DV_PAIRS=46
  run $1 $2 AVT_DA_SPLITS=9
done
echo "== cfg2 step"; python scripts/dev_stage_times.py 128 2>&1 | head -1
} > $T 2>&1
cat $T

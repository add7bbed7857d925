#!/bin/bash
# dev: A/B timings on one B200 (one process per env-switch variant)
mkdir -p gpurun_out
T=gpurun_out/ab_$1.txt
run() { local b=$1 na=$2; shift 2; env "$@" python scripts/dev_bwd_time.py $b $na 2>&1 | tail -1; }
{
echo "== un-profiled backward, B=1024"
run 1024 1024 AVT_X=base
run 1024 1024 AVT_DV_STAGGER_NS=4000
run 1024 1024 AVT_DV_STAGGER_NS=9000
run 1024 1024 AVT_DV_STAGGER_NS=14000
run 1024 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
run 1024 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1 AVT_DV_STAGGER_NS=6000
run 1024 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1 AVT_DV_STAGGER_NS=12000
run 1024 1024 AVT_DV_PAIRS=45 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
run 1024 1024 AVT_DV_PAIRS=40 AVT_DA_SPLITS=8 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
echo "== 512 rows x 1024 audio (N=2 shard)"
run 512 1024 AVT_X=base
run 512 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
run 512 1024 AVT_DV_STAGGER_NS=9000
echo "== 256 rows x 1024 audio (N=4 shard)"
run 256 1024 AVT_X=base
run 256 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
echo "== 128 rows x 1024 audio (N=8 shard)"
run 128 1024 AVT_X=base
run 128 1024 AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1
run 128 1024 AVT_DV_BMN=1
echo "== parity of the switched paths (gpu tests of the head with the switches on)"
AVT_DV_PAIRS=46 AVT_DA_SPLITS=7 AVT_DA_RR=1 AVT_BWD_REVERSE=1 AVT_DV_BMN=1 AVT_DV_STAGGER_NS=6000 timeout 600 python -m pytest tests/test_head_gpu.py -m gpu -x -q 2>&1 | tail -4
} > $T 2>&1
cat $T

#!/bin/bash
# dev: A/B timings on one B200 (stage times of the B=1024 step per variant, the dv cycle trace, shard times)
mkdir -p gpurun_out
{
echo "== default (AVT_DV_GROUPS=2)"; python scripts/dev_stage_times.py 2>&1 | tail -9
echo "== AVT_DV_GROUPS=4"; AVT_DV_GROUPS=4 python scripts/dev_stage_times.py 2>&1 | tail -9
echo "== trace build"; AVT_LIB_PATH=$PWD/build_variants/libavtubes_trace.so python scripts/dev_tc_trace.py 2>&1 | tail -12
echo "== trace build, AVT_DV_GROUPS=4"; AVT_DV_GROUPS=4 AVT_LIB_PATH=$PWD/build_variants/libavtubes_trace.so python scripts/dev_tc_trace.py 2>&1 | tail -4
echo "== shard times"; python scripts/dev_shard_times.py 2>&1 | tail -5
} > gpurun_out/ab_$1.txt 2>&1
cat gpurun_out/ab_$1.txt

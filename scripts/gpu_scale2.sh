#!/bin/bash
# 8-GPU box: NCCL parity tests, then bench.py at N = 2 and 8 with the collective extras (train clips/s, sharded eval)
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_dist_gpu.py -m gpu -x -q > gpurun_out/sd_pytest_dist.txt 2>&1; echo "dist tests rc=$?" | tee -a gpurun_out/sd_pytest_dist.txt
for N in 2 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) \
      bench.py --gpus $N --no-cpu-baseline > gpurun_out/sd_n$N.json 2> gpurun_out/sd_n$N.err; echo "n$N rc=$?"
done
tail -3 gpurun_out/sd_pytest_dist.txt
grep -o "world [0-9] rank 0 [^}]*}" gpurun_out/sd_pytest_dist.txt | cut -c1-700

#!/bin/bash
# one 8-GPU box: NCCL parity tests of the sharded step, then bench.py at N = 1, 2, 4, 8 back to back
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/sc_topo.txt 2>&1
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -x -q > gpurun_out/sc_pytest_dist.txt 2>&1; echo "dist tests rc=$?" | tee -a gpurun_out/sc_pytest_dist.txt
timeout 300 python bench.py --gpus 1 --no-extras --no-cpu-baseline > gpurun_out/sc_n1.json 2> gpurun_out/sc_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
      bench.py --gpus $N --no-cpu-baseline > gpurun_out/sc_n$N.json 2> gpurun_out/sc_n$N.err; echo "n$N rc=$?"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 \
      bench.py --gpus 8 --workload cfg2 --no-extras --no-cpu-baseline > gpurun_out/sc_cfg2_n8.json 2> gpurun_out/sc_cfg2_n8.err; echo "cfg2 n8 rc=$?"
tail -3 gpurun_out/sc_pytest_dist.txt
for N in 1 2 4 8; do python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sc_n$N.json").read().strip().splitlines()[-1])
    print($N, round(d["ms_per_step"],4), "%.3e"%d["value"], "loss", d["loss"], "e2e ms", round(d["e2e"]["ms_per_step"],3))
except Exception as e: print($N, "no line", e)
PY
done

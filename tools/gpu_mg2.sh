#!/bin/bash
# 2-GPU visit: slab parity worker (peer-pull and NCCL paths, point gather) and a short 2-GPU bench line.
O=gpurun_out/${1:-mg2}; mkdir -p $O
timeout 300 python -m pytest tests/test_multigpu.py -m gpu -q -s --timeout 200 > $O/pytest_mg.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "dist_worker|passed|failed|skipped" $O/pytest_mg.log | tail -6
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 \
  bench.py --gpus 2 --steps 60 --warmup 5 > $O/bench2.json 2> $O/bench2.err; echo "bench rc=$?" | tee -a $O/rc.txt
tail -1 $O/bench2.json | cut -c1-700

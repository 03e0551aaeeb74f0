#!/bin/bash
# Multi-GPU visit (gpurun --gpus 8): NCCL parity at 2/4/8 ranks, then the scaling sweep.
TAG=${1:-scale}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
for N in 1 2 4 8; do
  if [ $N = 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 60 --warmup 5 --no-extras > $O/bench_$N.json 2> $O/bench_$N.err
  else
    NCCL_DEBUG=WARN timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600+N)) bench.py --gpus $N --steps 60 --warmup 5 > $O/bench_$N.json 2> $O/bench_$N.err
  fi
  echo "bench $N rc=$?" | tee -a $O/rc.txt
done
tail -3 $O/pytest.log
for N in 1 2 4 8; do python - $O/bench_$N.json <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]; d=json.loads(l)
    print(d['n_gpus'], round(d['value'],1), 'Gvox/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value'],1), 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],3), d['clocks'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done

#!/bin/bash
# Short multi-GPU visit: bench at the rank counts given (default "8 4").
TAG=${1:-scale8}; shift
O=gpurun_out/$TAG; mkdir -p $O
for N in ${@:-8 4}; do
  NCCL_DEBUG=WARN timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29600+N)) bench.py --gpus $N --steps 200 --warmup 10 > $O/bench_$N.json 2> $O/bench_$N.err
  echo "bench $N rc=$?"
  python - $O/bench_$N.json <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]; d=json.loads(l)
    print(d['n_gpus'], round(d['value'],1), 'Gvox/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value'],1), 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],3), d['clocks'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done

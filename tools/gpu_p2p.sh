#!/bin/bash
# peer-pull vs ncclBroadcast on N GPUs (default 2): parity worker in both modes, then the bench
N=${1:-2}; O=gpurun_out/p2p_$N; mkdir -p $O
timeout 400 python -m pytest tests/test_multigpu.py -m gpu -x -q -s --timeout 300 ${PYTEST_K:+-k "$PYTEST_K"} 2>&1 | grep -E "dist_worker|passed|failed|Error|error" | tail -8
for mode in auto nccl; do
  [ $mode = nccl ] && export TSDF_B200_BCAST=nccl || unset TSDF_B200_BCAST
  NCCL_DEBUG=WARN timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29700+N)) bench.py --gpus $N --steps 200 --warmup 10 > $O/bench_$mode.json 2> $O/bench_$mode.err
  echo "bench $mode rc=$?"
  python - $O/bench_$mode.json <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]; d=json.loads(l)
    print(d['n_gpus'], d['config']['frame_distribution'][:12], round(d['value'],1), 'Gvox/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value'],1), 'kernel_ms', round(d['roofline']['kernel_ms'],4))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
  tail -2 $O/bench_$mode.err
done

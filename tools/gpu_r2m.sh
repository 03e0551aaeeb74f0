#!/bin/bash
# Round-2 visit E (8 GPUs): multi-rank parity at 2/4/8 ranks (the only place it can run), group tests across
# devices, bench at N = 8/4/2/1 as the driver runs it (+ per-frame timelines at 8), reference arm under torchrun,
# single-process group throughput.
O=gpurun_out/r2m; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 700 python -m pytest tests/test_multigpu.py -m gpu -v --timeout 300 > $O/pytest_multigpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "PASSED|FAILED|SKIPPED|passed|failed|dist_worker" $O/pytest_multigpu.log | tail -30
bench() { # N tag env...
  N=$1; tag=$2; shift; shift
  env "$@" NCCL_DEBUG=WARN timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29800+N)) bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_$tag.json 2> $O/bench_$tag.err
  echo "bench $tag rc=$?" | tee -a $O/rc.txt
  python - $O/bench_$tag.json <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]; d=json.loads(l)
    print(d['n_gpus'], d['frame_distribution'][:12], round(d['value'],1), 'Gvox/s', round(d['ms_per_step'],4), 'ms', [round(x,4) for x in d['windows']['ms_per_step']],
          'e2e', round(d['e2e']['value'],1), 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'launches', d['gpu_launches'], 'rb', round(d['e2e']['readback']['ms'],3), d['e2e']['readback']['points'], d['clocks'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
}
bench 8 n8 X=1
bench 4 n4 X=1
bench 2 n2 X=1
timeout 300 python bench.py --steps 20 --warmup 5 --no-extras > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2m/bench_n1.json') if l.startswith('{')][-1])
print('n1', round(d['value'],1), d['ms_per_step'], 'kernel', d['roofline']['kernel_ms'], 'frac', d['roofline']['frac'], d['clocks'])
PY
for n in 8; do timeout 300 python tools/group_bench.py $n 40 2>&1 | tail -1 | tee -a $O/group_bench.jsonl; done

#!/bin/bash
# Round-2 visit B (2 GPUs): whole GPU suite (multi-GPU + group tests run across devices), N=2 bench,
# and the N=2 emulation of the 8-rank slab load (DZ=256: 128 slices per rank) with the per-frame timeline.
O=gpurun_out/r2b; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1800 python -m pytest tests -m gpu -x -q --timeout 600 --durations=10 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -22 $O/pytest.log
run() { # tag, env..., then bench args
  tag=$1; shift
  env "$@" NCCL_DEBUG=WARN timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port 29712 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_$tag.json 2> $O/bench_$tag.err
  echo "bench $tag rc=$?" | tee -a $O/rc.txt
  python - $O/bench_$tag.json <<'PY'
import json,sys
try:
    l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]; d=json.loads(l)
    print(d['n_gpus'], d['frame_distribution'][:12], round(d['value'],1), 'Gvox/s', round(d['ms_per_step'],4), 'ms', [round(x,4) for x in d['windows']['ms_per_step']],
          'e2e', round(d['e2e']['value'],1), 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'launches', d['gpu_launches'], d['e2e']['gpu_launches'], 'rb', d['e2e']['readback']['ms'], d['e2e']['readback']['points'], d['clocks'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
  tail -2 $O/bench_$tag.err
}
run n2 X=1
run n2_slab8 TSDF_BENCH_DZ=256 TSDF_BENCH_TRACE=r2b
run n2_slab8_nccl TSDF_BENCH_DZ=256 TSDF_B200_BCAST=nccl
timeout 300 python bench.py --steps 20 --warmup 5 --no-extras > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b/bench_n1.json') if l.startswith('{')][-1])
print('n1', round(d['value'],1), d['ms_per_step'], d['windows']['ms_per_step'], 'cov', d['coverage'], 'frac', d['roofline']['frac'], 'kernel', d['roofline']['kernel_ms'], d['clocks'])
print('e2e', d['e2e']['value'], d['e2e']['readback'])
PY

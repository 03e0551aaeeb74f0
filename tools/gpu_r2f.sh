#!/bin/bash
# Round-2 visit F (2 GPUs): multi-rank parity worker (new: 10 frames, interleaved plain integrate, pinned
# gather, gathered mesh), group + C++ device-list tests across devices, reference arm under torchrun,
# sparse-kernel work counters.
O=gpurun_out/r2f; mkdir -p $O
timeout 900 python -m pytest tests/test_multigpu.py tests/test_group_gpu.py tests/test_cpp_host.py tests/test_abi.py tests/test_cull_gpu.py tests/test_exactmath_gpu.py -m gpu -v --timeout 300 > $O/pytest_multigpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
grep -E "PASSED|FAILED|SKIPPED|passed|failed|dist_worker|Error" $O/pytest_multigpu.log | tail -40
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29833 \
  bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $O/bench_ref_torchrun.json 2> $O/bench_ref_torchrun.err; echo "ref rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2f/bench_ref_torchrun.json') if l.startswith('{')][-1])
print('ref under torchrun', d['value'], d['ms_per_step'], d['cpu_baseline']['cores'], d['cpu_baseline']['omp_num_threads_env'])
PY
for s in real kitti; do timeout 200 python tools/sparse_counters.py $s 2>&1 | tail -1 | tee -a $O/sparse_counters.jsonl; done
for n in 2; do timeout 300 python tools/group_bench.py $n 40 2>&1 | tail -1 | tee -a $O/group_bench.jsonl; done

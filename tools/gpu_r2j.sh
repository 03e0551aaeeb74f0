#!/bin/bash
# Round-2 visit J (1 GPU): host staging copy A/B, run-safe A/B, suite, both bench arms.
O=gpurun_out/r2j; mkdir -p $O
for t in 3 7 0; do TSDF_B200_COPY_THREADS=$t timeout 120 python tools/hostcopy_ab.py 2>&1 | tail -1 | tee -a $O/hostcopy_ab.jsonl; done
L=$PWD/semantic-slam_b200
for rep in 1 2 3; do
 for a in "full 1024" "slab 1024"; do
  AB_ROUNDS=4 timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/runsafe   /' | tee -a $O/runsafe_ab.jsonl
  AB_ROUNDS=4 TSDF_B200_DEV_LIB=$L/libtsdf_b200_dev_norunsafe.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/norunsafe /' | tee -a $O/runsafe_ab.jsonl
 done
done
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest.log
for t in 3 0; do
TSDF_B200_COPY_THREADS=$t timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_t$t.json 2> $O/bench_t$t.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - $O/bench_t$t.json <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[1], 'value', round(d['value'],1), round(d['ms_per_step'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])
print(' multi', {k:round(v['ms_per_frame'],3) for k,v in d['extra_multi_frame_call'].items() if k.startswith('k=')})
print(' c2 fps', round(d['extra_c2']['frames_per_sec']), 'k', round(d['extra_c2']['kernel_ms'],4), ' c3 fps', round(d['extra_c3']['frames_per_sec']), 'k', round(d['extra_c3']['kernel_ms'],4), ' c5 kf/s', round(d['extra_c5']['frames_per_sec']), 'k', round(d['extra_c5']['kernel_ms'],4))
print(' refgpu', {k:(round(v['this_library_blocking_call_ms'],3), round(v['this_library_kernel_ms'],3)) for k,v in d['reference_gpu_kernel'].items() if k!='what'})
PY
done

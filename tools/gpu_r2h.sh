#!/bin/bash
# Round-2 visit H (1 GPU): suite, sparse A/B (quarter lists with / without prefetch-ahead vs plain), pair A/B, bench.
O=gpurun_out/r2h; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest.log
for a in "real 512" "kitti 1024" "real 256"; do
  timeout 300 python tools/variants.py $a 0 4194304 2097152 2>&1 | tail -1 | tee -a $O/sparse_ab.jsonl
done
for a in "full 1024" "slab 1024"; do
  AB_ROUNDS=5 timeout 300 python tools/pair_ab.py $a 2>&1 | tail -1 | tee -a $O/pair_ab.jsonl
done
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2h/bench.json') if l.startswith('{')][-1])
for k in ('value','ms_per_step','windows','coverage','gpu_launches','clocks'): print(k, d.get(k))
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('includes',)})
print('roofline', {k:v for k,v in d['roofline'].items() if k in ('achieved','frac','kernel_ms','algorithmic_bytes_per_launch')})
print('multi', d.get('extra_multi_frame_call'))
print('c2', d.get('extra_c2')); print('c3', d.get('extra_c3')); print('c5', d.get('extra_c5')); print('refgpu', d.get('reference_gpu_kernel'))
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_launch.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_integrate -s 3 -c 1 -f -o $O/integrate_full python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt

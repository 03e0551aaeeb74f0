#!/bin/bash
# Round-2 visit A (1 GPU): whole GPU suite incl. the new full-size parity tests, both bench arms.
O=gpurun_out/r2a; mkdir -p $O
nproc > $O/nproc.txt; free -g > $O/free.txt
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 --durations=15 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -25 $O/pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
tail -c 1500 $O/bench_ref.json; tail -3 $O/bench.err $O/bench_ref.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2a/bench.json') if l.startswith('{')][-1])
for k in ('value','ms_per_step','windows','coverage','gpu_launches','clocks'): print(k, d.get(k))
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('includes',)})
print('roofline', {k:v for k,v in d['roofline'].items() if k in ('achieved','frac','kernel_ms','algorithmic_bytes_per_launch')})
print('cpu', d.get('cpu_baseline'))
print('multi', d.get('extra_multi_frame_call'))
print('c2', d.get('extra_c2')); print('c3', d.get('extra_c3'))
PY

#!/bin/bash
# Round-2 visit I (1 GPU): suite (batched quarter kernel), bench with extras.
O=gpurun_out/r2i; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt; tail -2 $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2i/bench.json') if l.startswith('{')][-1])
for k in ('value','ms_per_step','windows','coverage','gpu_launches','clocks'): print(k, d.get(k))
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('includes','readback')}, d['e2e']['readback']['ms'])
print('roofline', {k:v for k,v in d['roofline'].items() if k in ('achieved','frac','kernel_ms','traffic')})
print('multi', d.get('extra_multi_frame_call'))
print('c2', d.get('extra_c2')); print('c3', d.get('extra_c3')); print('c5', d.get('extra_c5'))
PY

#!/bin/bash
# Round-2 visit C (1 GPU): whole GPU suite, pair-kernel A/B, bench with extras, ncu of the pair kernel.
O=gpurun_out/r2c; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -8 $O/pytest.log
for a in "full 1024" "sphere 1024" "slab 1024" "full 512"; do
  timeout 200 python tools/pair_ab.py $a 2>&1 | tail -1 | tee -a $O/pair_ab.jsonl
done
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2c/bench.json') if l.startswith('{')][-1])
for k in ('value','ms_per_step','windows','coverage','gpu_launches','clocks'): print(k, d.get(k))
print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('includes',)})
print('roofline', {k:v for k,v in d['roofline'].items() if k in ('achieved','frac','kernel_ms','algorithmic_bytes_per_launch')})
print('multi', d.get('extra_multi_frame_call'))
print('c2', d.get('extra_c2')); print('c3', d.get('extra_c3')); print('c5', d.get('extra_c5'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_integrate2 -s 2 -c 1 -f -o $O/integrate2_full \
   python tools/pair_ab.py full 1024 8 > $O/ncu_full.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt
ncu -i $O/integrate2_full.ncu-rep --page raw --csv > $O/integrate2_raw.csv 2>/dev/null

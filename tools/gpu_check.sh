#!/bin/bash
# One GPU-box visit: parity tests, smoke, both bench arms, ncu launch list, one full capture.
# usage (from the repo root on the box): bash tools/gpu_check.sh [tag]
TAG=${1:-run}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 400 python bench.py --impl reference --steps 10 --warmup 2 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
   python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_launch.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_integrate -s 3 -c 2 -f -o $O/integrate_full \
   python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest.log; tail -2 $O/smoke.log; cat $O/bench.json; cat $O/bench_ref.json

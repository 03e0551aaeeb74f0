#!/bin/bash
# One GPU-box visit centred on the mesh readback: its parity tests first (own log), memcheck of one
# of them, then the whole GPU suite, smoke and the default bench line.
# usage (from the repo root on the box): bash tools/gpu_mesh.sh [tag]
TAG=${1:-mesh}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/smi.txt 2>&1
timeout 200 python -m pytest tests/test_mesh.py -m gpu -q --timeout 120 > $O/pytest_mesh.log 2>&1; echo "mesh rc=$?" | tee -a $O/rc.txt
tail -25 $O/pytest_mesh.log
timeout 150 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_mesh.py -m gpu -q -k "noise and dims0" \
   > $O/memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a $O/rc.txt
tail -5 $O/memcheck.log
timeout 420 python -m pytest tests -m gpu -x -q --timeout 300 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -5 $O/pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
tail -3 $O/smoke.log
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
cat $O/bench.json

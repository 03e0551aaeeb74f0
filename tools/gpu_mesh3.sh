#!/bin/bash
# Short GPU-box visit: mesh + extraction parity (incl. the many-tile scan test), then the sweep A/B.
TAG=${1:-mesh}
O=gpurun_out/$TAG
mkdir -p $O
timeout 300 python -m pytest tests/test_mesh.py tests/test_parity_gpu.py -m gpu -q --timeout 200  > $O/pytest_mesh.log 2>&1; echo "mesh rc=$?" | tee -a $O/rc.txt
tail -8 $O/pytest_mesh.log
for a in "real 512" "sphere 1024"; do
  timeout 200 python tools/mesh_ab.py $a >> $O/mesh_ab.jsonl 2>> $O/mesh_ab.err; echo "ab($a) rc=$?" | tee -a $O/rc.txt
done
cat $O/mesh_ab.jsonl
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/mesh_launches.csv \
   env AB_ROUNDS=1 python tools/mesh_ab.py sphere 1024 > $O/ncu_list.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt
grep -c k_ $O/mesh_launches.csv

#!/bin/bash
# Short GPU-box visit: mesh parity tests, the sweep A/B, and an ncu capture of the mesh kernels.
TAG=${1:-mesh2}
O=gpurun_out/$TAG
mkdir -p $O
timeout 200 python -m pytest tests/test_mesh.py tests/test_cpp_host.py -m gpu -q --timeout 120 > $O/pytest_mesh.log 2>&1; echo "mesh rc=$?" | tee -a $O/rc.txt
tail -15 $O/pytest_mesh.log
for a in "real 512" "real 512 cu" "sphere 1024"; do
  timeout 200 python tools/mesh_ab.py $a >> $O/mesh_ab.jsonl 2>> $O/mesh_ab.err; echo "ab($a) rc=$?" | tee -a $O/rc.txt
done
cat $O/mesh_ab.jsonl
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mc_ -c 6 -f -o $O/mesh_full \
   env AB_ROUNDS=1 python tools/mesh_ab.py real 512 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
ncu -i $O/mesh_full.ncu-rep --page raw --csv > $O/mesh_full_raw.csv 2>/dev/null

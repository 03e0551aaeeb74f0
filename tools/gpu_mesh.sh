#!/bin/bash
# One GPU-box visit centred on the mesh readback: parity tests, the sweep A/B
# (8-wide / 4-wide / scalar, tools/mesh_ab.py), the ncu launch list of a 1024^3 run and, with
# FULL=1, one `ncu --set full` capture of the mesh kernels on the 512^3 scene.
# usage (from the repo root on the box): [FULL=1] bash tools/gpu_mesh.sh [tag]
TAG=${1:-mesh}
O=gpurun_out/$TAG
mkdir -p $O
timeout 300 python -m pytest tests/test_mesh.py tests/test_parity_gpu.py tests/test_cpp_host.py -m gpu -q --timeout 200 \
   > $O/pytest_mesh.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -5 $O/pytest_mesh.log
for a in "real 512" "real 512 cu" "sphere 1024"; do
  timeout 200 python tools/mesh_ab.py $a >> $O/mesh_ab.jsonl 2>> $O/mesh_ab.err; echo "ab($a) rc=$?" | tee -a $O/rc.txt
done
cat $O/mesh_ab.jsonl
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/mesh_launches.csv \
   env AB_ROUNDS=1 python tools/mesh_ab.py sphere 1024 > $O/ncu_list.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt
if [ -n "$FULL" ]; then
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mc_ -c 6 -f -o $O/mesh_full \
     env AB_ROUNDS=1 python tools/mesh_ab.py real 512 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
  ncu -i $O/mesh_full.ncu-rep --page raw --csv > $O/mesh_full_raw.csv 2>/dev/null
fi

#!/bin/bash
# Round-2 visit L (1 GPU): running-pointer row loop vs the previous build (dense + sparse), then the suite.
O=gpurun_out/r2l; mkdir -p $O
L=$PWD/semantic-slam_b200
for rep in 1 2 3; do
 for a in "full 1024" "slab 1024" "real 512" "kitti 1024"; do
  AB_ROUNDS=4 timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/new  /' | tee -a $O/ptr_ab.jsonl
  AB_ROUNDS=4 TSDF_B200_DEV_LIB=$L/libtsdf_b200_dev_prev.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/prev /' | tee -a $O/ptr_ab.jsonl
 done
done
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -4 $O/pytest.log

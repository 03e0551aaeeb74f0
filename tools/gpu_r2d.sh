#!/bin/bash
# Round-2 visit D (1 GPU): suite + A/Bs: free-space short cuts (bit 15), register cap (alt lib), pair kernel,
# staged image tile (bit 16, TMA) on the sparse configs; ncu of the staged kernel.
O=gpurun_out/r2d; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -6 $O/pytest.log
for rep in 1 2; do
  AB_ROUNDS=5 timeout 300 python tools/variants.py full 1024 0 32768 2>&1 | tail -1 | sed 's/^/mb4 /' | tee -a $O/free_ab.jsonl
  AB_ROUNDS=5 TSDF_B200_LIB=$PWD/semantic-slam_b200/libtsdf_b200_mb3.so timeout 300 python tools/variants.py full 1024 0 32768 2>&1 | tail -1 | sed 's/^/mb3 /' | tee -a $O/free_ab.jsonl
done
AB_ROUNDS=5 timeout 300 python tools/variants.py slab 1024 0 32768 2>&1 | tail -1 | sed 's/^/mb4 /' | tee -a $O/free_ab.jsonl
AB_ROUNDS=5 TSDF_B200_LIB=$PWD/semantic-slam_b200/libtsdf_b200_mb3.so timeout 300 python tools/variants.py slab 1024 0 32768 2>&1 | tail -1 | sed 's/^/mb3 /' | tee -a $O/free_ab.jsonl
for a in "full 1024" "slab 1024" "sphere 1024"; do
  AB_ROUNDS=5 timeout 300 python tools/pair_ab.py $a 2>&1 | tail -1 | tee -a $O/pair_ab.jsonl
done
for a in "real 512" "kitti 1024" "real 256"; do
  timeout 300 python tools/variants.py $a 0 65536 2>&1 | tail -1 | tee -a $O/stage_ab.jsonl
done
for a in "real 512" "kitti 1024"; do
  set -- $a
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_integrate -s 20 -c 2 -f -o $O/stage_$1 \
     env AB_ROUNDS=1 python tools/variants.py $a 65536 0 > $O/ncu_$1.log 2>&1; echo "ncu $1 rc=$?" | tee -a $O/rc.txt
done

#!/bin/bash
# Round-2 visit G (1 GPU): suite (new quarter work-list sparse kernel), dense-kernel A/B across builds
# (update_run4 with / without free-space short cuts vs the round-1 update), sparse A/B, counters.
O=gpurun_out/r2g; mkdir -p $O
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
tail -6 $O/pytest.log
L=$PWD/semantic-slam_b200
for rep in 1 2 3; do
 for a in "full 1024" "slab 1024"; do
  AB_ROUNDS=4 timeout 300 python tools/variants.py $a 0 32768 2>&1 | tail -1 | sed 's/^/run4   /' | tee -a $O/dense_ab.jsonl
  AB_ROUNDS=4 TSDF_B200_DEV_LIB=$L/libtsdf_b200_dev_legacy.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/legacy /' | tee -a $O/dense_ab.jsonl
  AB_ROUNDS=4 TSDF_B200_DEV_LIB=$L/libtsdf_b200_dev_legacy_mb1.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/legmb1 /' | tee -a $O/dense_ab.jsonl
 done
done
for a in "real 512" "kitti 1024" "real 256"; do
  timeout 300 python tools/variants.py $a 0 2097152 2>&1 | tail -1 | tee -a $O/sparse_ab.jsonl
done
for s in real kitti; do timeout 200 python tools/sparse_counters.py $s 2>&1 | tail -1 | tee -a $O/sparse_counters.jsonl; done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_integrate_q -s 20 -c 1 -f -o $O/sparse_q_real \
     env AB_ROUNDS=1 python tools/variants.py real 512 0 > $O/ncu_real.log 2>&1; echo "ncu rc=$?" | tee -a $O/rc.txt

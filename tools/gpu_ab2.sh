#!/bin/bash
O=gpurun_out/${1:-ab}; mkdir -p $O
for a in "full 1024" "full 512" "real 512" "kitti 1024"; do
  timeout 300 python tools/variants.py $a 64 128 0 2>&1 | tail -1 | tee -a $O/variants.jsonl
done

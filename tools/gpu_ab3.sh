#!/bin/bash
O=gpurun_out/${1:-ab}; mkdir -p $O
for a in "full 512" "real 512" "kitti 1024"; do
  timeout 300 python tools/variants.py $a 0 128 64 2>&1 | tail -1 | tee -a $O/variants.jsonl
done
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 2>&1 | tail -3

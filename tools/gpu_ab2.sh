#!/bin/bash
O=gpurun_out/${1:-ab}; mkdir -p $O
V=${2:-"1024 0 128"}
for a in "full 1024" "full 512" "slab 1024"; do
  timeout 300 python tools/variants.py $a $V 2>&1 | tail -1 | tee -a $O/variants.jsonl
done

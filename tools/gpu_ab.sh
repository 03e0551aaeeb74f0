#!/bin/bash
# Round-robin A/B of kernel variants on one GPU box (see tools/variants.py for the variant bits).
# usage: bash tools/gpu_ab.sh <tag> "<variants>" ["<scene> <dim>" ...]   then GPU parity tests
O=gpurun_out/${1:-ab}; mkdir -p $O
V=${2:-"0 128 64"}
shift; shift
[ $# -eq 0 ] && set -- "full 512" "real 512" "kitti 1024"
for a in "$@"; do
  timeout 300 python tools/variants.py $a $V 2>&1 | tail -1 | tee -a $O/variants.jsonl
done
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 2>&1 | tail -3

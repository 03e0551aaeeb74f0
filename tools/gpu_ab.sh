#!/bin/bash
# A/B on one box: previous build (tools/ab/libtsdf_prev.so) vs the current one, interleaved.
O=gpurun_out/${1:-ab}; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for rep in 1 2; do
  for a in "full 1024" "full 512"; do
    [ -f tools/ab/libtsdf_prev.so ] && TSDF_B200_LIB=$PWD/tools/ab/libtsdf_prev.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/prev /' | tee -a $O/variants.jsonl
    timeout 300 python tools/variants.py $a 0 128 64 2>&1 | tail -1 | sed 's/^/cur  /' | tee -a $O/variants.jsonl
  done
done
for a in "real 512" "kitti 1024"; do
  [ -f tools/ab/libtsdf_prev.so ] && TSDF_B200_LIB=$PWD/tools/ab/libtsdf_prev.so timeout 300 python tools/variants.py $a 0 2>&1 | tail -1 | sed 's/^/prev /' | tee -a $O/variants.jsonl
  timeout 300 python tools/variants.py $a 0 128 64 2>&1 | tail -1 | sed 's/^/cur  /' | tee -a $O/variants.jsonl
done

#!/bin/bash
# A/B of two builds of the library (separate processes, alternating): $1 = other .so, rest = variants.py args
OTHER=$1; shift
for rep in 1 2 3; do
  AB_ROUNDS=4 timeout 300 python tools/variants.py "$@" 2>&1 | tail -1 | sed 's/^/base  /' | cut -c1-260
  AB_ROUNDS=4 TSDF_B200_LIB=$PWD/$OTHER timeout 300 python tools/variants.py "$@" 2>&1 | tail -1 | sed 's/^/other /' | cut -c1-260
done

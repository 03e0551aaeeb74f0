#!/bin/bash
# One GPU-box visit, parameterised by stage names.
# usage (from the repo root on the box): bash tools/gpu_visit.sh <tag> <stage> [<stage> ...]
#   noalloc   tests/test_zz_no_alloc.py alone, in a fresh process
#   tests     the whole -m gpu suite (ends with test_zz_no_alloc.py)
#   smoke     __graft_entry__.smoke()
#   bench     both bench arms (product, reference)
#   list      ncu launch list of the short bench (gpu__time_duration per launch)
#   full      ncu --set full capture of the dense integrate kernel (whole grid)
#   slab      the same on a 1024x1024x128 grid (the slab one of 8 GPUs owns: 128-voxel rows)
#   mgpu      tests/test_multigpu.py + tests/test_group_gpu.py (needs >= 2 GPUs)
#   scale:N   headline bench at N ranks        c3:N / c5:N   bench.py --workload c3|c5 at N ranks
TAG=$1; shift
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
for stage in "$@"; do
  case $stage in
    noalloc) timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q --timeout 300 > $O/noalloc.log 2>&1; echo "noalloc rc=$?" | tee -a $O/rc.txt; tail -5 $O/noalloc.log ;;
    tests) timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -5 $O/pytest.log ;;
    smoke) timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt; tail -2 $O/smoke.log ;;
    bench) timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt; cat $O/bench.json
           timeout 400 python bench.py --impl reference --steps 10 --warmup 2 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt; cat $O/bench_ref.json ;;
    list) timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
            python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_launch.log 2>&1; echo "ncu-list rc=$?" | tee -a $O/rc.txt ;;
    full) timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_integrate -s 3 -c 2 -f -o $O/integrate_full \
            python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt ;;
    slab) TSDF_BENCH_DZ=128 TSDF_BENCH_WINDOWS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_integrate -s 3 -c 2 -f -o $O/integrate_slab \
            python bench.py --steps 6 --warmup 3 --no-extras > $O/ncu_slab.log 2>&1; echo "ncu-slab rc=$?" | tee -a $O/rc.txt ;;
    mgpu) timeout 900 python -m pytest tests/test_multigpu.py tests/test_group_gpu.py -m gpu -q -s --timeout 300 > $O/mgpu.log 2>&1; echo "mgpu rc=$?" | tee -a $O/rc.txt
          grep -E "dist_worker|passed|failed|rror" $O/mgpu.log | tail -12 ;;
    scale:*|c3:*|c5:*)
          W=${stage%%:*}; N=${stage##*:}; [ $W = scale ] && WL="" || WL="--workload $W"
          if [ $N = 1 ]; then LAUNCH="python"; else LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N))"; fi
          NCCL_DEBUG=WARN timeout 400 $LAUNCH bench.py --gpus $N --steps 200 --warmup 10 $WL --no-extras > $O/${W}_$N.json 2> $O/${W}_$N.err
          echo "$W N=$N rc=$?" | tee -a $O/rc.txt; grep '^{' $O/${W}_$N.json | tail -1 | cut -c1-1500; tail -2 $O/${W}_$N.err ;;
    *) echo "unknown stage $stage" ;;
  esac
done

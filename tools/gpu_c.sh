bash tools/gpu_check.sh c1 2>&1 | tail -12 | cut -c1-1500
bash tools/gpu_ab2.sh c1 | cut -c1-500

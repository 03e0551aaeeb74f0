bash tools/gpu_scale8.sh s3_cta4 8
TSDF_B200_NCCL_MAX_CTAS=1 bash tools/gpu_scale8.sh s3_cta1 8
TSDF_B200_NCCL_MAX_CTAS=0 bash tools/gpu_scale8.sh s3_cta0 8
timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -2

"""Multi-GPU parity (z-slabs + NCCL frame broadcast + point gather) — needs >= 2 GPUs."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,mode", [(2, "auto"), (2, "nccl"), (4, "auto"), (8, "auto")])
def test_rank_slabs_match_oracle(pkg, world, mode):
    """mode auto: peer-pull frame distribution when cudaIpc works on the box (the worker prints
    which one ran), nccl: the ncclBroadcast path forced."""
    if pkg.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ)
    if mode == "nccl":
        env["TSDF_B200_BCAST"] = "nccl"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29517 + world + (50 if mode == "nccl" else 0)),
           os.path.join(ROOT, "tests", "dist_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert f"dist_worker ok {world}" in r.stdout
    print(r.stdout.strip().splitlines()[-1])

"""GPU (needs >= 2 devices, else skipped): the data-parallel update -- NCCL all-reduce and the peer-memory exchange,
fp32 and bf16 -- equals a single-GPU update on the concatenated batch, and all ranks end with bit-identical weights."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
def test_n_gpu_step_equals_single_gpu_step_on_concatenated_batch(world):
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip("needs %d CUDA devices" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
           "--master-addr", "127.0.0.1", "--master-port", str(29611 + world), os.path.join(ROOT, "tests", "dp_identity_script.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0 and "DP_IDENTITY_OK" in r.stdout, r.stdout[-4000:]

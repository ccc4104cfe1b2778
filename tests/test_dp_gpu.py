"""GPU (needs >= 2 devices, else skipped): the NCCL data-parallel step equals a single-GPU step on the
concatenated batch, and all ranks end with bit-identical weights."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_two_gpu_step_equals_single_gpu_step_on_concatenated_batch():
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", os.path.join(ROOT, "tests", "dp_identity_script.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "DP_IDENTITY_OK" in r.stdout, r.stdout[-3000:]

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


def test_multi_gpu_launcher_trains_on_two_gpus():
    """scripts/dqn_multi_gpu.sh (the reference's command line) -> torchrun -> main.py on 2 GPUs: episodes of different
    length per rank (different seeds), peer-memory exchange inside the captured update, every rank ends normally."""
    import tempfile
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    logdir = tempfile.mkdtemp(prefix="dqn_dp_main_")
    env = dict(os.environ, DQN_LOG_ROOT=logdir, DQN_MASTER_PORT="29655", DQN_DP_TIMEOUT_S="60",
               DQN_EXTRA_FLAGS="--num_episodes 3 --max_steps_per_episode 80 --replay_memory_capacity 2000 "
                               "--replay_start_size 64 --summary_freq 4 --seed 3")
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "dqn_multi_gpu.sh"), "atari", "SyntheticPong-v0", "2"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-4000:]
    assert "global_step/sec" in r.stdout and "Episode = 2" in r.stdout, r.stdout[-4000:]

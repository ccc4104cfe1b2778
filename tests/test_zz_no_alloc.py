"""GPU: the allocation-free forward entry point against the handle-based one.

``dqn_cnn_forward_noalloc`` exists for hosts that are not allowed to allocate device memory (a
process confined to a pre-assigned slice of the GPU, a latency-critical loop that must stay out of
the allocator, a sandboxed plugin).  It keeps its intermediate activations in an array built into
the library, shared by all calls -- so this test calls it twice in a row with different inputs
and batch sizes and checks each result against ``dqn_net_forward`` (whose buffers come from
cudaMalloc); the second call also proves that leftovers of the first do not leak into it.
Collected last on purpose (file name): a failure here cannot hide other tests under ``-x``.
"""
import numpy as np
import pytest

from gpu_helpers import make_config, oracle_params, rel_err

pytestmark = pytest.mark.gpu


def test_noalloc_forward_matches_handle_based_forward_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import _cabi as c
    import network
    A = 6
    ishape = (84, 84, 4)
    net = network.Network.create_network(make_config(network="cnn", minibatch_size=64, max_batch=64), ishape, A)
    params = oracle_params("cnn", ishape, A, 21)
    net.set_params(params)
    dev = net.device
    flat = torch.as_tensor(np.concatenate([np.asarray(p, np.float32).ravel() for p in params]), device=dev)
    assert flat.numel() == 143776 + 257 * A
    rng = np.random.default_rng(4)
    for n in (64, 5):                                   # two calls in a row; the second is smaller
        x = torch.as_tensor(rng.integers(0, 256, (n,) + ishape, dtype=np.uint8), device=dev)
        q_ref, am_ref = net.predict(x, argmax=True)     # existing entry point: dqn_net_forward
        q = torch.full((n, A), float("nan"), dtype=torch.float32, device=dev)
        am = torch.full((n,), -1, dtype=torch.int32, device=dev)
        c.check(c.lib.dqn_cnn_forward_noalloc(flat.data_ptr(), A, x.data_ptr(), n, q.data_ptr(), am.data_ptr(),
                                              c.current_stream()), "noalloc forward")
        torch.cuda.synchronize()
        assert rel_err(q.cpu().numpy(), q_ref.cpu().numpy()) <= 1e-4      # the tolerance of the fp32 parity tests
        assert torch.equal(am, am_ref)
    # beyond the built-in scratch: refused with a clear message, nothing launched
    x = torch.zeros((65,) + ishape, dtype=torch.uint8, device=dev)
    q = torch.zeros((65, A), dtype=torch.float32, device=dev)
    rc = c.lib.dqn_cnn_forward_noalloc(flat.data_ptr(), A, x.data_ptr(), 65, q.data_ptr(), None, c.current_stream())
    assert rc == -1 and "at most 64 states" in c.last_error()

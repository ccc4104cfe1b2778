import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/dist-dqn_b200'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
from gpu_helpers import *
import network
from oracle import net_ref as N
ARCH, ISHAPE, A = "cnn", (84,84,4), 6
for B in (5, 70):
    params = oracle_params(ARCH, ISHAPE, A, 3)
    s, a, r, s2, t = make_batch(ISHAPE, A, B, 4)
    t[:] = 1
    r = np.random.default_rng(8).uniform(-2000.0, 2000.0, B)
    nets = {}
    grads = {}
    for prec in ("fp32", "bf16"):
        cfg = make_config(network=ARCH, minibatch_size=B, max_batch=max(B, 8), precision=prec)
        net = network.Network.create_network(cfg, ISHAPE, A)
        net.set_params(params); net.sync_target()
        net.backward_on_batch(to_device_batch(net, s, a, r, s2, t), include_reg=False)
        grads[prec] = net.get_grads()
    print("B", B, "bf16 vs fp32 (data term):", ["%.4f" % rel_err(g, e) for g, e in zip(grads["bf16"], grads["fp32"])])
    # correlation / scale of each tensor
    for i, (g, e) in enumerate(zip(grads["bf16"], grads["fp32"])):
        g = g.ravel().astype(np.float64); e = e.ravel().astype(np.float64)
        print(i, "scale", float((g @ e) / (e @ e + 1e-300)), "cos", float((g @ e) / (np.linalg.norm(g) * np.linalg.norm(e) + 1e-300)))

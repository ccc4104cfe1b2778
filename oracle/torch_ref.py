"""torch-CPU autograd restatement of the reference's TF graph (TEST INFRASTRUCTURE).

Independent of ``net_ref.py`` (library conv/pool + autograd instead of hand-written im2col
and backward).  Two uses: (a) cross-check of ``net_ref.py`` in float64; (b) the timed
"reference-on-CPU (restated)" baseline of ``bench.py`` in float32 on all host cores,
driven by a transcription of ``DQNAgent._train_minibatch``'s three-call structure
(``dqn_agent.py:108-143``).  PARITY UNPINNED (see ``oracle/__init__.py``).
"""
from __future__ import annotations

import math
import torch
import torch.nn.functional as F

from .net_ref import ARCHS, conv_plan, reg_indices


def forward(arch, params, x, input_shape):
    """params: list of torch tensors in the reference tuple order, TF layouts (HWIO, [in,out])."""
    spec = ARCHS[arch]
    dt = params[0].dtype
    x = x.to(dt)
    if spec["kind"] == "mlp":
        w1, b1, w2, b2, w3, b3 = params                       # network.py:303-308
        a1 = torch.tanh(x @ w1 + b1)
        a2 = torch.tanh(a1 @ w2 + b2)
        return a2 @ w3 + b3
    layers, flat = conv_plan(arch, input_shape)
    a = x.permute(0, 3, 1, 2)                                  # NHWC -> NCHW for torch
    for i, L in enumerate(layers):                             # network.py:416-424
        w = params[2 * i].permute(3, 2, 0, 1)                  # HWIO -> OIHW
        a = F.pad(a, (L["pad_w"][0], L["pad_w"][1], L["pad_h"][0], L["pad_h"][1]))
        a = torch.relu(F.conv2d(a, w, params[2 * i + 1], stride=L["s"]))
        if L["pool"]:
            a = F.pad(a, (0, 2 * L["pw"] - L["ow"], 0, 2 * L["ph"] - L["oh"]), value=float("-inf"))
            a = F.max_pool2d(a, 2, 2)
    n = len(layers)
    w4, b4, w5, b5 = params[2 * n:2 * n + 4]
    a3 = a.permute(0, 2, 3, 1).reshape(a.shape[0], flat)       # flatten in NHWC order (network.py:401)
    a4 = torch.relu(a3 @ w4 + b4)
    return a4 @ w5 + b5                                        # network.py:404


def loss_fn(arch, params, x, y, actions, reg_param, input_shape):
    q = forward(arch, params, x, input_shape)
    onehot = F.one_hot(actions.long(), q.shape[1]).to(q.dtype)
    q_masked = (q * onehot).sum(1)                             # network.py:149
    loss = ((q_masked - y.to(q.dtype)) ** 2).mean()            # network.py:150
    reg = sum(0.5 * (params[i] ** 2).sum() for i in reg_indices(arch))
    return loss + reg_param * reg, q                           # network.py:151-152


def loss_and_grads(arch, params, x, y, actions, reg_param, input_shape):
    ps = [p.detach().clone().requires_grad_(True) for p in params]
    loss, q = loss_fn(arch, ps, x, y, actions, reg_param, input_shape)
    grads = torch.autograd.grad(loss, ps)
    return loss.detach(), q.detach(), [g.detach() for g in grads]


class CpuBaselineAgent:
    """Transcription of the reference step for CPU timing: RMSProp/Adam/SGD only (the presets)."""

    def __init__(self, arch, input_shape, num_actions, optimizer, lr, reg_param, gamma,
                 rmsprop_decay=0.95, target_update_freq=10000, seed=0, dtype=torch.float32):
        from .net_ref import init_params
        self.arch, self.input_shape, self.A = arch, tuple(input_shape), num_actions
        self.opt, self.lr, self.reg, self.gamma, self.decay = optimizer, lr, reg_param, gamma, rmsprop_decay
        self.tfreq = target_update_freq
        self.params = [torch.tensor(p, dtype=dtype) for p in init_params(arch, input_shape, num_actions, seed)]
        self.target = [p.clone() for p in self.params]
        self.s0 = [torch.ones_like(p) if optimizer == "rmsprop" else torch.zeros_like(p) for p in self.params]
        self.s1 = [torch.zeros_like(p) for p in self.params]
        self.steps = 0

    @torch.no_grad()
    def predict(self, states_list, use_target=False):
        import numpy as np
        x = torch.from_numpy(np.asarray(states_list, dtype=np.float32))   # the feed conversion (dqn_agent.py:186-189)
        return forward(self.arch, self.target if use_target else self.params, x, self.input_shape).numpy()

    def train_minibatch(self, replay, k):
        """dqn_agent.py:108-143 with ``replay`` an oracle ReplayRef holding real stacks."""
        import numpy as np
        from .replay_ref import ReplayRef, one_hot_ref
        nt, t = replay.get_minibatch(k)
        nt, t = list(nt), list(t)
        nxt = list(ReplayRef.get_next_states(nt))
        maxq = self.predict(nxt, True).max(axis=1) if nxt else np.zeros((0,), np.float32)
        states, yq, acts = [], [], []
        tq = list(maxq) + [0] * len(t)                          # zip_longest(fillvalue=0), dqn_agent.py:242
        for item, target_q in zip(nt + t, tq):
            s, a, r, _, _ = item
            states.append(s)
            yq.append(float(r) + self.gamma * float(target_q))  # python float64 arithmetic
            acts.append(one_hot_ref(a, self.A))
        x = torch.from_numpy(np.asarray(states, dtype=np.float32))
        y = torch.tensor(np.asarray(yq, dtype=np.float32))
        a = torch.from_numpy(np.asarray(acts, dtype=np.float32))
        ps = [p.requires_grad_(True) for p in self.params]
        q = forward(self.arch, ps, x, self.input_shape)
        loss = (((q * a).sum(1) - y) ** 2).mean()
        loss = loss + self.reg * sum(0.5 * (ps[i] ** 2).sum() for i in reg_indices(self.arch))
        grads = torch.autograd.grad(loss, ps)
        self.steps += 1
        with torch.no_grad():
            for p, g, s0, s1 in zip(self.params, grads, self.s0, self.s1):
                p.requires_grad_(False)
                if self.opt == "rmsprop":
                    s0.add_((g * g - s0) * (1.0 - self.decay))
                    p.sub_(self.lr * g / torch.sqrt(s0 + 1e-10))
                elif self.opt == "adam":
                    t_ = self.steps
                    lr_t = self.lr * math.sqrt(1 - 0.999 ** t_) / (1 - 0.9 ** t_)
                    s0.add_((g - s0) * 0.1)
                    s1.add_((g * g - s1) * 0.001)
                    p.sub_(lr_t * s0 / (torch.sqrt(s1) + 1e-8))
                else:
                    p.sub_(self.lr * g)
            if self.steps % self.tfreq == 0:
                self.target = [p.clone() for p in self.params]
        return float(loss)

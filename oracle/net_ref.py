"""numpy float64 restatement of the reference's TensorFlow graph (TEST INFRASTRUCTURE).

PARITY UNPINNED: TensorFlow is not available here and the reference has no golden
vectors, so this file *defines* the semantics the CUDA path is checked against.  It
follows, line by line:

* ``network.py:59-116``  graph assembly (placeholders, params, layers, loss, optimizer, target)
* ``network.py:141-157`` loss  = mean_b (sum_a q*onehot - y)^2 + reg_param * reg_loss
* ``network.py:159-203`` optimizer table (7 TF optimizers, ``learning_rate=config.lr``)
* ``network.py:258-313`` SimpleNetwork  (tanh MLP in->20->20->A, L2 on w1,w2,w3)
* ``network.py:317-424`` ConvNetwork    (3 x [conv SAME + bias + relu + 2x2/2 SAME max-pool],
                                          fc 256 relu, affine head, L2 on w4,w5)
* ``dqn_agent.py:108-143, 224-253`` TD target y = r + gamma * max_a Q_target(s', a), 0 for terminals

TF r0.9-era semantics restated here:
 (1) conv2d NHWC/HWIO is cross-correlation; SAME: out=ceil(in/s), pad_total=max((out-1)s+k-in,0),
     before=pad_total//2, rest after.  VALID: out=(in-k)//s+1, no pad.
 (2) max_pool SAME uses the same pad rule; padded cells never win; the gradient goes to the
     first maximum in (row, col) scan order of the window.
 (3) relu'(0)=0.  (4) l2_loss(w)=sum(w^2)/2.  (5) reduce_mean over the batch.
 (6) RMSProp: ms0=1, mom0=0; ms+=(1-decay)(g^2-ms); mom=momentum*mom+lr*g/sqrt(ms+eps); w-=mom;
     momentum=0, eps=1e-10 by default.
 (7) Adam: m=v=0; lr_t=lr*sqrt(1-b2^t)/(1-b1^t); m+=(1-b1)(g-m); v+=(1-b2)(g^2-v);
     w-=lr_t*m/(sqrt(v)+eps); b1=.9 b2=.999 eps=1e-8.
 (8) SGD w-=lr*g.  Momentum: acc=mu*acc+g; w-=lr*acc.
 (9) Adagrad: acc0=0.1; acc+=g^2; w-=lr*g/sqrt(acc).
(10) Adadelta: rho=.95 eps=1e-8; acc=rho*acc+(1-rho)g^2; u=sqrt(acc_u+eps)/sqrt(acc+eps)*g;
     acc_u=rho*acc_u+(1-rho)u^2; w-=lr*u.
(11) Ftrl: lr_power=-0.5, acc0=0.1, l1=l2=0; new=acc+g^2; lin+=g-(sqrt(new)-sqrt(acc))/lr*w;
     w=-lin/(sqrt(new)/lr); acc=new.

The "nature" architecture (VALID convs 8x8/4, 4x4/2, 3x3/1, no pooling, fc 512) is the
north_star's throughput variant; it is not in the reference.
"""
from __future__ import annotations

import math
import numpy as np

# ----------------------------------------------------------------------------- architectures
# conv entry: (kernel, stride, out_channels, padding, pool)
ARCHS = {
    # network.py:318-333, 389-424
    "cnn": dict(kind="conv",
                convs=[(8, 4, 32, "SAME", True), (4, 2, 64, "SAME", True), (3, 1, 64, "SAME", True)],
                fc=256),
    # Mnih et al. 2015 (north_star variant, not in the reference)
    "nature": dict(kind="conv",
                   convs=[(8, 4, 32, "VALID", False), (4, 2, 64, "VALID", False), (3, 1, 64, "VALID", False)],
                   fc=512),
    # network.py:258-313
    "simple": dict(kind="mlp", hidden=(20, 20)),
}


def out_and_pad(n_in: int, k: int, s: int, padding: str):
    """TF output size and (before, after) padding along one spatial axis."""
    if padding == "SAME":
        out = -(-n_in // s)
        total = max((out - 1) * s + k - n_in, 0)
        return out, total // 2, total - total // 2
    if padding == "VALID":
        return (n_in - k) // s + 1, 0, 0
    raise ValueError(padding)


def conv_plan(arch: str, input_shape):
    """Per conv layer: dict(in_h,in_w,cin,k,s,cout,pad_h,pad_w,oh,ow,pool,ph,pw). Returns (layers, flat)."""
    spec = ARCHS[arch]
    h, w, c = input_shape
    layers = []
    for (k, s, cout, padding, pool) in spec["convs"]:
        oh, pbh, pah = out_and_pad(h, k, s, padding)
        ow, pbw, paw = out_and_pad(w, k, s, padding)
        L = dict(in_h=h, in_w=w, cin=c, k=k, s=s, cout=cout, pad_h=(pbh, pah), pad_w=(pbw, paw),
                 oh=oh, ow=ow, pool=pool)
        if pool:
            L["ph"], _, _ = out_and_pad(oh, 2, 2, "SAME")
            L["pw"], _, _ = out_and_pad(ow, 2, 2, "SAME")
        else:
            L["ph"], L["pw"] = oh, ow
        layers.append(L)
        h, w, c = L["ph"], L["pw"], cout
    return layers, h * w * c


def param_shapes(arch: str, input_shape, num_actions: int):
    """Shapes in the reference's tuple order (network.py:296, 387): w1,b1,w2,b2,..."""
    spec = ARCHS[arch]
    if spec["kind"] == "mlp":
        if len(input_shape) != 1:
            raise RuntimeError("SimpleNetwork expects 1-d input")     # network.py:265-266
        h1, h2 = spec["hidden"]
        return [(input_shape[0], h1), (h1,), (h1, h2), (h2,), (h2, num_actions), (num_actions,)]
    if len(input_shape) != 3:
        raise RuntimeError("ConvNetwork expects 3-d input")            # network.py:338-339
    layers, flat = conv_plan(arch, input_shape)
    shapes = []
    for L in layers:
        shapes += [(L["k"], L["k"], L["cin"], L["cout"]), (L["cout"],)]
    fc = spec["fc"]
    # NB the reference hard-codes w4 as [256,256] (network.py:372-373); equal to [flat,fc] for 84x84x4.
    shapes += [(flat, fc), (fc,), (fc, num_actions), (num_actions,)]
    return shapes


def reg_indices(arch: str):
    """Indices (into the param tuple) of the weights under L2 (network.py:311, 407)."""
    return [0, 2, 4] if ARCHS[arch]["kind"] == "mlp" else [6, 8]


def init_params(arch, input_shape, num_actions, seed=0, stddev=0.01, dtype=np.float64):
    """truncated_normal(stddev) weights (redraw beyond 2 sigma), zero biases (network.py:269-270, 341-342)."""
    rng = np.random.default_rng(seed)
    out = []
    for shp in param_shapes(arch, input_shape, num_actions):
        if len(shp) == 1:
            out.append(np.zeros(shp, dtype))
        else:
            v = rng.standard_normal(shp)
            bad = np.abs(v) > 2.0
            while bad.any():
                v[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(v) > 2.0
            out.append((v * stddev).astype(dtype))
    return out


# ----------------------------------------------------------------------------- primitive ops
def _im2col(xp, k, s, oh, ow):
    B, H, W, C = xp.shape
    sb, sh, sw, sc = xp.strides
    v = np.lib.stride_tricks.as_strided(xp, (B, oh, ow, k, k, C), (sb, sh * s, sw * s, sh, sw, sc),
                                        writeable=False)
    return v.reshape(B * oh * ow, k * k * C)


def conv_fwd(x, w, L):
    xp = np.pad(x, ((0, 0), L["pad_h"], L["pad_w"], (0, 0)))
    cols = _im2col(xp, L["k"], L["s"], L["oh"], L["ow"])
    y = cols @ w.reshape(-1, L["cout"])
    return y.reshape(x.shape[0], L["oh"], L["ow"], L["cout"]), cols


def conv_bwd(dy, cols, w, L, need_dx=True):
    B = dy.shape[0]
    k, s, oh, ow, cin = L["k"], L["s"], L["oh"], L["ow"], L["cin"]
    dy2 = dy.reshape(-1, L["cout"])
    dw = (cols.T @ dy2).reshape(w.shape)
    if not need_dx:
        return dw, None
    dcols = (dy2 @ w.reshape(-1, L["cout"]).T).reshape(B, oh, ow, k, k, cin)
    hp = L["in_h"] + sum(L["pad_h"])
    wp = L["in_w"] + sum(L["pad_w"])
    dxp = np.zeros((B, hp, wp, cin), dy.dtype)
    for ky in range(k):
        for kx in range(k):
            dxp[:, ky:ky + s * oh:s, kx:kx + s * ow:s, :] += dcols[:, :, :, ky, kx, :]
    dx = dxp[:, L["pad_h"][0]:L["pad_h"][0] + L["in_h"], L["pad_w"][0]:L["pad_w"][0] + L["in_w"], :]
    return dw, dx


def pool_fwd(a, L):
    """2x2/2 SAME max pool; returns pooled and argmax in {0,1,2,3} = dy*2+dx (first max wins)."""
    B, oh, ow, C = a.shape
    ph, pw = L["ph"], L["pw"]
    ap = np.full((B, ph * 2, pw * 2, C), -np.inf, a.dtype)
    ap[:, :oh, :ow, :] = a                       # SAME pool pad is (0, after) for 2x2/2
    win = ap.reshape(B, ph, 2, pw, 2, C).transpose(0, 1, 3, 2, 4, 5).reshape(B, ph, pw, 4, C)
    arg = win.argmax(axis=3)                     # numpy argmax returns the first maximum
    return np.take_along_axis(win, arg[:, :, :, None, :], 3)[:, :, :, 0, :], arg.astype(np.uint8)


def pool_bwd(dp, arg, L):
    B, ph, pw, C = dp.shape
    dwin = np.zeros((B, ph, pw, 4, C), dp.dtype)
    np.put_along_axis(dwin, arg[:, :, :, None, :].astype(np.int64), dp[:, :, :, None, :], 3)
    dap = dwin.reshape(B, ph, pw, 2, 2, C).transpose(0, 1, 3, 2, 4, 5).reshape(B, ph * 2, pw * 2, C)
    return dap[:, :L["oh"], :L["ow"], :]


# ----------------------------------------------------------------------------- bf16 emulation
def bf16_round(a):
    """Round to the nearest bfloat16 (ties to even), returned in the input's float type.  Used to restate
    where the bf16 tensor-core mode of the CUDA path rounds: GEMM operands (weights, stored activations,
    propagated gradients); accumulation, bias, ReLU, pooling and the head stay in full precision."""
    a = np.asarray(a)
    f = a.astype(np.float32)
    u = f.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).astype(a.dtype if a.dtype.kind == "f" else np.float64)


def f16_round(a):
    """Round to the nearest IEEE half (ties to even), returned in the input's float type."""
    a = np.asarray(a)
    return a.astype(np.float16).astype(a.dtype if a.dtype.kind == "f" else np.float64)


def _ident(a):
    return a


# conv1 of the reference net ("cnn") in the CUDA path's 16-bit mode: f16 operands (the default kernels) or bf16 (the
# DQN_NO_RESIDENT=1 kernels).  Tests that exercise the latter set this to False.
CONV1_F16 = True


# ----------------------------------------------------------------------------- forward / backward
def forward(arch, params, x, input_shape=None, keep=False, bf16=False):
    """Q-values [B, A] (network.py:303-308 / 389-404).  x: [B]+input_shape, any dtype (cast, no /255)."""
    spec = ARCHS[arch]
    dt = params[0].dtype
    x = np.asarray(x).astype(dt)
    cache = {}
    Q = bf16_round if bf16 else _ident
    if spec["kind"] == "mlp":
        w1, b1, w2, b2, w3, b3 = params
        a1 = np.tanh(x @ w1 + b1)
        a2 = np.tanh(a1 @ w2 + b2)
        q = a2 @ w3 + b3
        if keep:
            cache = dict(x=x, a1=a1, a2=a2)
        return (q, cache) if keep else q
    layers, flat = conv_plan(arch, x.shape[1:] if input_shape is None else input_shape)
    a = x
    per = []
    for i, L in enumerate(layers):
        w, b = Q(params[2 * i]), params[2 * i + 1]
        if bf16 and arch == "cnn" and i == 0 and CONV1_F16:
            # conv1 of the reference net runs with f16 operands in the CUDA path's 16-bit mode (the uint8 frames enter the
            # tensor core as f16 without a conversion instruction, and both operands of an MMA share one format)
            w = f16_round(params[0])
        z, cols = conv_fwd(a, w, L)
        r = np.maximum(z + b, 0.0)
        if L["pool"]:
            p, arg = pool_fwd(r, L)
        else:
            p, arg = r, None
        p = Q(p)                                  # stored activation (bf16 mode rounds here)
        if keep:
            per.append(dict(cols=cols, gate=(p > 0), arg=arg, L=L, wq=w))
        a = p
    n = len(layers)
    w4, b4, w5, b5 = params[2 * n:2 * n + 4]
    a3 = a.reshape(a.shape[0], flat)             # NHWC flatten: (y*W+x)*C+c  (network.py:401)
    w4q = Q(w4)
    a4 = np.maximum(a3 @ w4q + b4, 0.0)          # the head reads this in full precision in both modes
    q = a4 @ w5 + b5
    if keep:
        cache = dict(per=per, a3=a3, a4=a4, a4_gate=(Q(a4) > 0), w4q=w4q, pooled_shape=a.shape)
        return q, cache
    return q


def loss_and_grads(arch, params, x, y, actions, reg_param, input_shape=None, bf16=False):
    """loss (network.py:149-152) and dL/dparam in tuple order.  actions: int [B].
    bf16=True restates the rounding points of the CUDA path's bf16 tensor-core mode (see bf16_round)."""
    spec = ARCHS[arch]
    Q = bf16_round if bf16 else _ident
    q, c = forward(arch, params, x, input_shape, keep=True, bf16=bf16)
    B, A = q.shape
    actions = np.asarray(actions, np.int64)
    y = np.asarray(y).astype(q.dtype)
    onehot = np.zeros((B, A), q.dtype)
    onehot[np.arange(B), actions] = 1.0
    q_masked = (q * onehot).sum(1)
    diff = q_masked - y
    reg = sum(0.5 * np.sum(params[i] ** 2) for i in reg_indices(arch))
    loss = np.mean(diff ** 2) + reg_param * reg
    dq = (2.0 / B) * diff[:, None] * onehot
    grads = [None] * len(params)
    if spec["kind"] == "mlp":
        w1, b1, w2, b2, w3, b3 = params
        grads[4] = c["a2"].T @ dq
        grads[5] = dq.sum(0)
        d2 = (dq @ w3.T) * (1.0 - c["a2"] ** 2)
        grads[2] = c["a1"].T @ d2
        grads[3] = d2.sum(0)
        d1 = (d2 @ w2.T) * (1.0 - c["a1"] ** 2)
        grads[0] = c["x"].T @ d1
        grads[1] = d1.sum(0)
    else:
        n = len(c["per"])
        w4, b4, w5, b5 = params[2 * n:2 * n + 4]
        grads[2 * n + 2] = c["a4"].T @ dq
        grads[2 * n + 3] = dq.sum(0)
        d4 = Q(dq @ w5.T) * c["a4_gate"]
        grads[2 * n] = c["a3"].T @ d4
        grads[2 * n + 1] = d4.sum(0)
        da = Q(d4 @ c["w4q"].T).reshape(c["pooled_shape"])
        for i in range(n - 1, -1, -1):
            P = c["per"][i]
            L = P["L"]
            # gate on the POOLED activation: identical to relu'(r) at the arg-max position
            dz = pool_bwd(da * P["gate"], P["arg"], L) if L["pool"] else da * P["gate"]
            dw, dx = conv_bwd(dz, P["cols"], P["wq"], L, need_dx=(i > 0))
            grads[2 * i] = dw
            grads[2 * i + 1] = dz.sum((0, 1, 2))
            da = Q(dx) if dx is not None else None
    for i in reg_indices(arch):
        grads[i] = grads[i] + reg_param * params[i]
    return loss, q, grads


# ----------------------------------------------------------------------------- optimizers
OPTIMIZERS = ("sgd", "momentum", "rmsprop", "adam", "adagrad", "adadelta", "ftrl")   # network.py:169-177


def init_slots(opt, params):
    z = lambda: [np.zeros_like(p) for p in params]
    if opt == "sgd":
        return []
    if opt == "momentum":
        return [z()]
    if opt == "rmsprop":
        return [[np.ones_like(p) for p in params], z()]            # ms=1, mom=0
    if opt == "adam":
        return [z(), z()]
    if opt == "adagrad":
        return [[np.full_like(p, 0.1) for p in params]]
    if opt == "adadelta":
        return [z(), z()]
    if opt == "ftrl":
        return [[np.full_like(p, 0.1) for p in params], z()]       # accum, linear
    raise RuntimeError("Unsupported optimizer {}".format(opt))       # network.py:179-180


def apply_optimizer(opt, params, grads, slots, step, lr, momentum=0.9, rmsprop_decay=0.95):
    """In-place TF update; ``step`` is the 1-based count of this apply (Adam's t)."""
    for i, (w, g) in enumerate(zip(params, grads)):
        if opt == "sgd":
            w -= lr * g
        elif opt == "momentum":
            acc = slots[0][i]
            acc *= momentum
            acc += g
            w -= lr * acc
        elif opt == "rmsprop":
            ms, mom = slots[0][i], slots[1][i]
            ms += (1.0 - rmsprop_decay) * (g * g - ms)
            mom *= 0.0                                            # TF default momentum=0.0
            mom += lr * g / np.sqrt(ms + 1e-10)
            w -= mom
        elif opt == "adam":
            m, v = slots[0][i], slots[1][i]
            b1, b2, eps = 0.9, 0.999, 1e-8
            lr_t = lr * math.sqrt(1.0 - b2 ** step) / (1.0 - b1 ** step)
            m += (1.0 - b1) * (g - m)
            v += (1.0 - b2) * (g * g - v)
            w -= lr_t * m / (np.sqrt(v) + eps)
        elif opt == "adagrad":
            acc = slots[0][i]
            acc += g * g
            w -= lr * g / np.sqrt(acc)
        elif opt == "adadelta":
            acc, acc_u = slots[0][i], slots[1][i]
            rho, eps = 0.95, 1e-8
            acc *= rho
            acc += (1.0 - rho) * g * g
            u = np.sqrt(acc_u + eps) / np.sqrt(acc + eps) * g
            acc_u *= rho
            acc_u += (1.0 - rho) * u * u
            w -= lr * u
        elif opt == "ftrl":
            acc, lin = slots[0][i], slots[1][i]
            new = acc + g * g
            lin += g - (np.sqrt(new) - np.sqrt(acc)) / lr * w
            w[...] = -lin / (np.sqrt(new) / lr)
            acc[...] = new
        else:
            raise RuntimeError("Unsupported optimizer {}".format(opt))


# ----------------------------------------------------------------------------- TD target + step
def td_targets(rewards, terminals, q_next_target, gamma):
    """dqn_agent.py:118-122, 242-246: y = r + gamma*max_a Q_tgt(s') (float64), 0 bootstrap for terminals,
    then rounded to float32 by the q_placeholder (network.py:62)."""
    r = np.asarray(rewards, np.float64)
    mx = np.asarray(q_next_target).max(axis=1).astype(np.float64)
    y = r + float(gamma) * np.where(np.asarray(terminals, bool), 0.0, mx)
    return y.astype(np.float32)


class RefTrainer:
    """Holds online/target params + optimizer slots and applies the reference update (dqn_agent.py:108-143)."""

    def __init__(self, arch, input_shape, num_actions, optimizer="rmsprop", lr=0.00025, reg_param=0.001,
                 gamma=0.99, momentum=0.9, rmsprop_decay=0.95, target_update_freq=10000, seed=0,
                 dtype=np.float64, params=None, bf16=False):
        self.bf16 = bf16        # restate the rounding points of the CUDA bf16 tensor-core mode
        self.arch, self.input_shape, self.num_actions = arch, tuple(input_shape), num_actions
        self.opt, self.lr, self.reg, self.gamma = optimizer, lr, reg_param, gamma
        self.momentum, self.decay, self.tfreq = momentum, rmsprop_decay, target_update_freq
        self.params = [np.array(p, dtype) for p in params] if params is not None else \
            init_params(arch, input_shape, num_actions, seed, dtype=dtype)
        self.slots = init_slots(optimizer, self.params)
        self.target = [p.copy() for p in self.params]             # dqn_agent.py:50 (step 0 sync)
        self.training_steps = 0

    def q_values(self, x, use_target=False):
        return forward(self.arch, self.target if use_target else self.params, x, self.input_shape, bf16=self.bf16)

    def train(self, states, actions, rewards, next_states, terminals):
        y = td_targets(rewards, terminals, self.q_values(next_states, True), self.gamma)
        loss, q, grads = loss_and_grads(self.arch, self.params, states, y.astype(np.float64), actions,
                                        self.reg, self.input_shape, bf16=self.bf16)
        self.training_steps += 1
        apply_optimizer(self.opt, self.params, grads, self.slots, self.training_steps, self.lr,
                        self.momentum, self.decay)
        if self.training_steps % self.tfreq == 0:                  # dqn_agent.py:143, 220
            self.target = [p.copy() for p in self.params]
        return loss, q, grads, y


def flops_per_sample(arch, input_shape, num_actions):
    """(fwd MACs, train MACs) per sample: online fwd + target fwd + wgrad all + dgrad all but first layer."""
    spec = ARCHS[arch]
    if spec["kind"] == "mlp":
        h1, h2 = spec["hidden"]
        macs = [input_shape[0] * h1, h1 * h2, h2 * num_actions]
    else:
        layers, flat = conv_plan(arch, input_shape)
        macs = [L["oh"] * L["ow"] * L["k"] ** 2 * L["cin"] * L["cout"] for L in layers]
        macs += [flat * spec["fc"], spec["fc"] * num_actions]
    fwd = sum(macs)
    return fwd, 2 * fwd + fwd + (fwd - macs[0])

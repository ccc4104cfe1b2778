"""Network builders with the reference's surface (reference src/network.py:13-424), backed by
hand-written CUDA kernels instead of a TensorFlow graph.

``Network.create_network(config, input_shape, num_actions, ...)`` returns an object exposing the
nine attributes the reference's callers use (``x_placeholder``, ``q_placeholder``,
``action_placeholder``, ``q_output``, ``train_op``, ``target_q_output``, ``target_update_ops``,
``summary_op``, ``global_step``).  Here they are opaque handles that ``Session.run`` understands,
so code written against the reference (``session.run(network.train_op, feed_dict=...)``) still
works; the fast path is ``network.train_step(replay_memory, k)``: one sample-to-update call with
no host round trip (gather -> both forwards -> TD target -> loss -> backward -> optimizer).

Parameters are one flat fp32 device buffer per set, tensors in the reference's tuple order
(network.py:296, 387) and TF layouts; ``get_params`` / ``set_params`` move them as numpy arrays.
"""
from __future__ import annotations

import numpy as np

import _cabi as _c


class _Handle(object):
    """Stand-in for a TF tensor/op handle."""

    def __init__(self, network, name):
        self.network, self.name = network, name

    def __repr__(self):
        return "<dqn_b200 handle %s>" % self.name


class Network(object):
    x_placeholder = None
    q_placeholder = None
    action_placeholder = None

    q_output = None
    train_op = None

    target_q_output = None
    target_update_ops = None

    summary_op = None
    global_step = None

    ARCH = None          # set by subclasses

    def __init__(self, input_shape, num_actions, num_replicas=1, ps_device=None, worker_device=None):
        self.input_shape = list(input_shape)
        self.num_actions = num_actions
        self.num_replicas = num_replicas      # data-parallel world size (gradient mean, network.py:187-192)
        self.ps_device = ps_device            # accepted for compatibility; parameters are replicated per GPU
        self.worker_device = worker_device
        self._h = None

    # ------------------------------------------------------------------ factory (network.py:35-57)
    @staticmethod
    def create_network(config, input_shape, num_actions, num_replicas=1, ps_device=None, worker_device=None):
        Net = {
            'simple': SimpleNetwork,
            'cnn': ConvNetwork,
            'nature': NatureNetwork,
        }.get(config.network, None)
        if Net is None:
            raise RuntimeError('Unsupported network type {}'.format(config.network))
        net = Net(input_shape=input_shape, num_actions=num_actions, num_replicas=num_replicas,
                  ps_device=ps_device, worker_device=worker_device)
        net._init_network(config)
        return net

    # ------------------------------------------------------------------ construction
    def _check_input_shape(self):
        raise NotImplementedError

    def _init_network(self, config):
        import torch
        self._check_input_shape()
        opt = getattr(config, 'optimizer', 'sgd')
        if opt not in _c.OPTIMIZER:
            raise RuntimeError('Unsupported optimizer {}'.format(opt))          # network.py:179-180
        self.config = config
        dev_index = getattr(config, 'device_index', None)
        self.device = torch.device('cuda', torch.cuda.current_device() if dev_index is None else int(dev_index))
        self.max_batch = int(max(getattr(config, 'max_batch', 0) or 0, getattr(config, 'minibatch_size', 32), 32))
        self.in_dtype = getattr(config, 'input_dtype', None) or ('f64' if self.ARCH == 'simple' else 'u8')
        shape = (_c.c_i32 * len(self.input_shape))(*self.input_shape)
        h = _c.c_vp()
        _c.check(_c.lib.dqn_net_create(_c.C.byref(h), _c.ARCH[self.ARCH], shape, len(self.input_shape),
                                       int(self.num_actions), self.max_batch, _c.DTYPE[self.in_dtype],
                                       self.device.index), type(self).__name__)
        self._h = h
        n, nt = _c.c_i64(), _c.c_i32()
        _c.check(_c.lib.dqn_net_param_count(h, _c.C.byref(n), _c.C.byref(nt)))
        self.flat_size, self.num_tensors = n.value, nt.value
        self.tensors = []
        for i in range(nt.value):
            off, numel, rank = _c.c_i64(), _c.c_i64(), _c.c_i32()
            shp = (_c.c_i64 * 4)()
            _c.check(_c.lib.dqn_net_tensor_info(h, i, _c.C.byref(off), _c.C.byref(numel), _c.C.byref(rank), shp))
            self.tensors.append((off.value, numel.value, tuple(shp[:rank.value])))
        self.precision = getattr(config, 'precision', None) or 'fp32'
        if self.precision not in ('fp32', 'bf16'):
            raise RuntimeError('Unsupported precision {}'.format(self.precision))
        if self.precision == 'bf16':
            _c.check(_c.lib.dqn_net_set_precision(h, 1), type(self).__name__)
        self.set_hyper(config, reset_slots=True)
        seed = getattr(config, 'seed', None)
        self.set_params(self._initial_params(np.random.default_rng(seed)))
        self.sync_target()
        # opaque handles (network.py:61-64, 77, 93, 103)
        for name in ('x_placeholder', 'q_placeholder', 'action_placeholder', 'q_output', 'train_op',
                     'target_q_output', 'target_update_ops', 'summary_op', 'global_step'):
            setattr(self, name, _Handle(self, name))
        self._loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._io = {}
        self.peers_connected = False

    def __del__(self):
        try:
            if self._h is not None:
                _c.lib.dqn_net_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _initial_params(self, rng):
        """truncated_normal(stddev=0.01) weights, zero biases (network.py:269-270, 341-342)."""
        out = []
        for (_, numel, shape) in self.tensors:
            if len(shape) == 1:
                out.append(np.zeros(shape, np.float32))
            else:
                v = rng.standard_normal(shape)
                bad = np.abs(v) > 2.0
                while bad.any():
                    v[bad] = rng.standard_normal(int(bad.sum()))
                    bad = np.abs(v) > 2.0
                out.append((v * 0.01).astype(np.float32))
        return out

    # ------------------------------------------------------------------ hyper-parameters
    def set_hyper(self, config, reset_slots=False, grad_scale=None):
        hp = _c.Hyper()
        hp.gamma = float(getattr(config, 'reward_discount', 0.9))
        hp.lr = float(getattr(config, 'lr', 0.001))
        hp.reg_param = float(getattr(config, 'reg_param', 0.001))
        hp.momentum = float(getattr(config, 'momentum', 0.9))
        hp.rmsprop_decay = float(getattr(config, 'rmsprop_decay', 0.95))
        opt = getattr(config, 'optimizer', 'sgd')
        if opt not in _c.OPTIMIZER:
            raise RuntimeError('Unsupported optimizer {}'.format(opt))
        hp.optimizer = _c.OPTIMIZER[opt]
        hp.loss = _c.LOSS[getattr(config, 'loss', 'mse') or 'mse']
        hp.grad_scale = float(grad_scale if grad_scale is not None else 1.0)
        _c.check(_c.lib.dqn_net_set_hyper(self._h, _c.C.byref(hp), 1 if reset_slots else 0), 'Network')
        self._hyper = hp

    # ------------------------------------------------------------------ parameters
    def _flatten(self, tensors):
        flat = np.zeros(self.flat_size, np.float32)
        if len(tensors) != len(self.tensors):
            raise RuntimeError('expected %d tensors, got %d' % (len(self.tensors), len(tensors)))
        for (off, numel, shape), t in zip(self.tensors, tensors):
            t = np.asarray(t, np.float32)
            if t.shape != shape:
                raise RuntimeError('tensor shape %s does not match %s' % (t.shape, shape))
            flat[off:off + numel] = t.ravel()
        return flat

    def _unflatten(self, flat):
        return [flat[off:off + numel].reshape(shape).copy() for (off, numel, shape) in self.tensors]

    def set_params(self, tensors, which='online'):
        flat = self._flatten(tensors)
        _c.check(_c.lib.dqn_net_upload(self._h, _c.BUF[which], flat.ctypes.data, flat.size), 'set_params')

    def get_params(self, which='online'):
        flat = np.empty(self.flat_size, np.float32)
        _c.check(_c.lib.dqn_net_download(self._h, _c.BUF[which], flat.ctypes.data, flat.size), 'get_params')
        return self._unflatten(flat)

    def flat_buffer(self, which='grad'):
        """torch view (no copy) of one of the flat device buffers, e.g. for the gradient all-reduce."""
        import torch
        p, n = _c.c_vp(), _c.c_i64()
        _c.check(_c.lib.dqn_net_buffer(self._h, _c.BUF[which], _c.C.byref(p), _c.C.byref(n)))
        return _wrap_device_memory(p.value, n.value, self.device)

    # ------------------------------------------------------------------ compute
    def _state_tensor(self, states):
        """list/ndarray/tensor of states -> contiguous device tensor in the net's input dtype."""
        import torch
        tdt = {'u8': torch.uint8, 'f32': torch.float32, 'f64': torch.float64}[self.in_dtype]
        if isinstance(states, torch.Tensor):
            t = states
        else:
            ndt = {'u8': np.uint8, 'f32': np.float32, 'f64': np.float64}[self.in_dtype]
            arr = np.asarray(states)
            if arr.dtype != ndt:
                if self.in_dtype == 'u8' and not np.array_equal(arr, arr.astype(np.uint8)):
                    raise RuntimeError('this network was built for uint8 frames; build it with '
                                       'config.input_dtype="f32" to feed non-integer states')
                arr = arr.astype(ndt)
            t = torch.from_numpy(np.ascontiguousarray(arr))
        t = t.to(device=self.device, dtype=tdt).contiguous()
        if list(t.shape[1:]) != self.input_shape:
            raise RuntimeError('states have shape %s, expected [N]+%s' % (list(t.shape), self.input_shape))
        return t

    def predict(self, states, use_target_network=False, argmax=False):
        """DQNAgent._predict_q_values (dqn_agent.py:176-189) -> device tensor [N, A] (and greedy actions)."""
        import torch
        x = self._state_tensor(states)
        n = x.shape[0]
        q = torch.empty((n, self.num_actions), dtype=torch.float32, device=self.device)
        am = torch.empty(n, dtype=torch.int32, device=self.device) if argmax else None
        which = _c.BUF['target' if use_target_network else 'online']
        for s in range(0, n, self.max_batch):
            e = min(n, s + self.max_batch)
            _c.check(_c.lib.dqn_net_forward(self._h, which, x[s:e].data_ptr(), e - s, q[s:e].data_ptr(),
                                            am[s:e].data_ptr() if argmax else None, _c.current_stream()), 'predict')
        return (q, am) if argmax else q

    def train_on_batch(self, batch, want_loss=False, want_q=False):
        """One update from a device ``Batch`` (replay_memory.Batch)."""
        import torch
        k = batch.actions.numel()
        q = torch.empty((k, self.num_actions), dtype=torch.float32, device=self.device) if want_q else None
        _c.check(_c.lib.dqn_net_train_step(self._h, batch.states.data_ptr(), batch.next_states.data_ptr(),
                                           batch.actions.data_ptr(), batch.rewards.data_ptr(), batch.terminals.data_ptr(),
                                           k, self._loss.data_ptr() if want_loss else None,
                                           q.data_ptr() if want_q else None, _c.current_stream()), 'train_step')
        return (self._loss if want_loss else None), q

    def backward_on_batch(self, batch, want_loss=False, want_q=False, include_reg=False):
        """Gradients only (flat DQN_BUF_GRAD buffer): the data-parallel half-step, and parity checks."""
        import torch
        k = batch.actions.numel()
        q = torch.empty((k, self.num_actions), dtype=torch.float32, device=self.device) if want_q else None
        _c.check(_c.lib.dqn_net_backward(self._h, batch.states.data_ptr(), batch.next_states.data_ptr(),
                                         batch.actions.data_ptr(), batch.rewards.data_ptr(), batch.terminals.data_ptr(),
                                         k, self._loss.data_ptr() if want_loss else None,
                                         q.data_ptr() if want_q else None, 1 if include_reg else 0,
                                         _c.current_stream()), 'backward')
        return (self._loss if want_loss else None), q

    def apply_gradients(self):
        _c.check(_c.lib.dqn_net_apply(self._h, _c.current_stream()), 'apply')

    def get_grads(self):
        flat = np.empty(self.flat_size, np.float32)
        _c.check(_c.lib.dqn_net_download(self._h, _c.BUF['grad'], flat.ctypes.data, flat.size), 'get_grads')
        return self._unflatten(flat)

    def train_step(self, replay_memory, k, d_indices=None, want_loss=False, loss_out=None):
        """Sample-to-update in one C call: gather (ring -> stacks) + the whole update.  ``loss_out``: a 1-element
        float32 tensor the loss kernel writes to instead of the network's own device scalar (a pinned host tensor
        works: the device writes it over PCIe, which saves a copy node in a captured step)."""
        if d_indices is None:
            d_indices = replay_memory.upload_indices(replay_memory.sample_indices(k))
        out = (loss_out if loss_out is not None else self._loss) if want_loss else None
        _c.check(_c.lib.dqn_train_step(self._h, replay_memory.handle, d_indices.data_ptr(), int(k),
                                       out.data_ptr() if want_loss else None, _c.current_stream()), 'train_step')
        return out

    def train_step_table(self, replay_memory, index_table, want_loss=False, loss_out=None, data_parallel=False):
        """``train_step`` (or ``dp_train_step``) on the current row of a ``replay_memory.IndexTable``: the update reads
        its pre-drawn sample indices from the device-resident table and advances the table's cursor itself, so a
        captured update replays with no per-update work on the indices (dqn_train_step_table)."""
        out = (loss_out if loss_out is not None else self._loss) if want_loss else None
        _c.check(_c.lib.dqn_train_step_table(self._h, replay_memory.handle, _c.ptr(index_table.table), index_table.rows,
                                             index_table.k, _c.ptr(index_table.cursor), _c.ptr(index_table.scratch),
                                             out.data_ptr() if want_loss else None, 1 if data_parallel else 0,
                                             _c.current_stream()), 'train_step_table')
        return out

    # ------------------------------------------------------------------ checkpoint / resume
    def save(self, path, extra=None):
        """Everything needed to resume training -- the reference got this from tf.train.Supervisor's Saver
        (main.py:136-143, 155): online and target parameters, both optimizer slots, global_step -- in one
        .npz file.  ``extra`` (dict of small python values, e.g. epsilon / training_steps) rides along."""
        blob = {'arch': np.array(self.ARCH), 'num_actions': np.array(self.num_actions),
                'input_shape': np.array(self.input_shape), 'global_step': np.array(self.get_global_step())}
        for which in ('online', 'target', 'slot0', 'slot1'):
            for i, t in enumerate(self.get_params(which)):
                blob['%s_%d' % (which, i)] = t
        for k, v in (extra or {}).items():
            blob['extra_' + k] = np.array(v)
        np.savez(path, **blob)

    def load(self, path):
        """Restores a checkpoint written by ``save`` (same architecture / action count); returns the extras."""
        z = np.load(path, allow_pickle=False)
        if str(z['arch']) != self.ARCH or int(z['num_actions']) != self.num_actions:
            raise RuntimeError('checkpoint is for %s with %d actions' % (z['arch'], int(z['num_actions'])))
        for which in ('online', 'target', 'slot0', 'slot1'):
            self.set_params([z['%s_%d' % (which, i)] for i in range(self.num_tensors)], which=which)
        # a connected data-parallel net realigns its exchange flags with the restored step: every rank loads at the
        # same point, with no update in flight anywhere
        if self.peers_connected:
            import torch
            import torch.distributed as dist
            torch.cuda.synchronize()
            dist.barrier()
        _c.check(_c.lib.dqn_net_set_global_step(self._h, int(z['global_step'])), 'load')
        if self.peers_connected:
            dist.barrier()
        return {k[6:]: z[k].item() for k in z.files if k.startswith('extra_')}

    def refresh_operands(self, which='online'):
        """After an in-place write to ``flat_buffer('online'/'target')`` (e.g. a broadcast): rebuilds the bf16 operand
        copies the tensor-core kernels read.  No-op in fp32 mode."""
        _c.check(_c.lib.dqn_net_refresh_operands(self._h, _c.BUF[which], _c.current_stream()), 'refresh_operands')

    def broadcast_parameters(self, src=0):
        """Rank ``src`` initialises, everyone else copies (the reference's chief / non-chief start-up,
        main.py:133-143); the target network starts as a copy of the online one (dqn_agent.py:50)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.broadcast(self.flat_buffer('online'), src)
            self.refresh_operands('online')
        self.sync_target()

    # ------------------------------------------------------------------ data parallel over NVLink peer memory
    def connect_peers(self):
        """Maps every rank's gradient exchange block (CUDA IPC) so the optimizer kernel can sum the ranks'
        gradients itself; afterwards ``dp_train_step`` / ``dp_train_on_batch`` replace backward + all-reduce +
        apply.  Needs an initialised torch.distributed group with one rank per GPU of one box."""
        import torch
        import torch.distributed as dist
        world, rank = dist.get_world_size(), dist.get_rank()
        mine = (_c.C.c_uint8 * 64)()
        _c.check(_c.lib.dqn_net_dp_handle(self._h, mine), 'connect_peers')
        gathered = [None] * world
        dist.all_gather_object(gathered, bytes(mine))
        blob = b''.join(gathered)
        _c.check(_c.lib.dqn_net_dp_connect(self._h, rank, world, blob), 'connect_peers')
        self.set_hyper(self.config, grad_scale=1.0 / world)
        self.peers_connected = True
        torch.cuda.synchronize()
        dist.barrier()

    def dp_status(self):
        """Raises if a peer-exchange wait of this rank ever timed out (a peer stalled longer than the timeout set with
        ``DQN_DP_TIMEOUT_S`` / dqn_net_dp_set_timeout, default 120 s): the weights are then no longer in sync."""
        s = _c.c_i64()
        _c.check(_c.lib.dqn_net_dp_status(self._h, _c.C.byref(s), _c.current_stream()), 'dp_status')
        if s.value:
            raise RuntimeError('data-parallel exchange timed out waiting for a peer at update %d' % s.value)

    def dp_train_on_batch(self, batch, want_loss=False):
        k = batch.actions.numel()
        _c.check(_c.lib.dqn_net_dp_step(self._h, batch.states.data_ptr(), batch.next_states.data_ptr(),
                                        batch.actions.data_ptr(), batch.rewards.data_ptr(), batch.terminals.data_ptr(),
                                        k, self._loss.data_ptr() if want_loss else None, _c.current_stream()), 'dp_step')
        return self._loss if want_loss else None

    def dp_train_step(self, replay_memory, k, d_indices=None, want_loss=False, loss_out=None):
        if d_indices is None:
            d_indices = replay_memory.upload_indices(replay_memory.sample_indices(k))
        out = (loss_out if loss_out is not None else self._loss) if want_loss else None
        _c.check(_c.lib.dqn_train_step_dp(self._h, replay_memory.handle, d_indices.data_ptr(), int(k),
                                          out.data_ptr() if want_loss else None, _c.current_stream()),
                 'dp_train_step')
        return out

    def sync_target(self):
        """The assign ops of network.py:251-252 as one device memcpy."""
        _c.check(_c.lib.dqn_net_target_sync(self._h, _c.current_stream()), 'target_sync')

    def get_global_step(self):
        s = _c.c_i64()
        _c.check(_c.lib.dqn_net_global_step(self._h, _c.C.byref(s), _c.current_stream()))
        return s.value

    @property
    def handle(self):
        return self._h


def _wrap_device_memory(ptr, nfloats, device):
    """torch float32 tensor aliasing raw device memory (via __cuda_array_interface__)."""
    import torch

    class _Mem(object):
        pass
    m = _Mem()
    m.__cuda_array_interface__ = {'shape': (int(nfloats),), 'typestr': '<f4', 'data': (int(ptr), False), 'version': 2}
    return torch.as_tensor(m, device=device)


class SimpleNetwork(Network):
    """Two tanh hidden layers + affine head (network.py:258-313)."""
    HIDDEN1_SIZE = 20
    HIDDEN2_SIZE = 20
    ARCH = 'simple'

    def _check_input_shape(self):
        if len(self.input_shape) != 1:
            raise RuntimeError('%s expects 1-d input' % type(self).__name__)


class ConvNetwork(Network):
    """The reference's conv net (network.py:317-424): SAME convs, each followed by a 2x2 max-pool, fc 256."""
    CONV1_FILTERS = 32
    CONV1_SIZE = 8
    CONV1_STRIDE = 4

    CONV2_FILTERS = 64
    CONV2_SIZE = 4
    CONV2_STRIDE = 2

    CONV3_FILTERS = 64
    CONV3_SIZE = 3
    CONV3_STRIDE = 1

    POOL_SIZE = [1, 2, 2, 1]
    POOL_STRIDE = [1, 2, 2, 1]

    FULLY_CONNECTED_SIZE = 256
    ARCH = 'cnn'

    def _check_input_shape(self):
        if len(self.input_shape) != 3:
            raise RuntimeError('%s expects 3-d input' % type(self).__name__)


class NatureNetwork(ConvNetwork):
    """Mnih et al. 2015 layout (VALID convs, no pooling, fc 512) -- the north_star's throughput
    variant; not part of the reference."""
    FULLY_CONNECTED_SIZE = 512
    ARCH = 'nature'


class Session(object):
    """Minimal ``session.run(fetches, feed_dict)`` so code written against the reference's TF
    session (dqn_agent.py:132-138, 189, 222) keeps working.  Slow path: every call moves its
    feeds from the host; use ``Network.train_step`` for training throughput."""

    graph = None

    def run(self, fetches, feed_dict=None):
        single = not isinstance(fetches, (list, tuple))
        items = [fetches] if single else list(fetches)
        # target_update_ops is fetched as the handle itself (the reference passes the op list)
        feed = {k.name: v for k, v in (feed_dict or {}).items()}
        out = []
        trained = None
        for f in items:
            net, name = f.network, f.name
            if name in ('q_output', 'target_q_output'):
                q = net.predict(feed['x_placeholder'], use_target_network=(name == 'target_q_output'))
                out.append(q.cpu().numpy())
            elif name in ('train_op', 'summary_op'):
                if trained is None:
                    trained = self._train(net, feed)
                out.append(None if name == 'train_op' else trained)
            elif name == 'target_update_ops':
                net.sync_target()
                out.append(None)
            elif name == 'global_step':
                out.append(net.get_global_step())
            else:
                raise RuntimeError('cannot fetch %r' % (f,))
        return out[0] if single else out

    @staticmethod
    def _train(net, feed):
        """train_op with host-computed targets (dqn_agent.py:249-253): expected_q enters as the
        reward of a terminal transition, so y = expected_q exactly."""
        import torch
        from replay_memory import Batch
        x = net._state_tensor(feed['x_placeholder'])
        y = torch.as_tensor(np.asarray(feed['q_placeholder'], np.float32).astype(np.float64), device=net.device)
        acts = torch.as_tensor(np.asarray(feed['action_placeholder']).argmax(axis=1).astype(np.int32), device=net.device)
        term = torch.ones(y.numel(), dtype=torch.uint8, device=net.device)
        if x.shape[0] > net.max_batch:
            raise RuntimeError('minibatch %d exceeds max_batch %d' % (x.shape[0], net.max_batch))
        loss, _ = net.train_on_batch(Batch(x, x, acts, y, term, None), want_loss=True)
        return float(loss.item())

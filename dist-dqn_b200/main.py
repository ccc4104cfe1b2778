#!/usr/bin/env python3
"""Entry point with the reference's flags (reference src/main.py:12-98) driving the B200 path.

The reference's parameter-server/gRPC cluster (main.py:100-193) is replaced by synchronous data
parallelism: launch one process per GPU with torchrun (scripts/dqn_multi_gpu.sh) and every rank trains
on its own replay shard while the gradients are averaged -- by default inside the step's own kernels over
NVLink peer memory (--dp peer), or with an NCCL all-reduce (--dp nccl); both have the reference's --sync
semantics.  The distribution flags are still accepted so the scripts keep working; --ps_hosts/
--worker_hosts/--job/--disable_* have no effect here.

--precision picks the arithmetic of the conv nets: bf16 (tcgen05 tensor cores, fp32 accumulation and
fp32 master weights; the default for cnn / nature on uint8 frames) or fp32 (CUDA cores, the 1e-4 parity
mode; always used by the simple network).
"""
import argparse
import logging
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument('--log_level', default='INFO', help='Log verbosity')
    # Environment
    p.add_argument('--env', default='CartPole-v0', help='Environment name')
    p.add_argument('--monitor', action='store_true', help='Accepted for compatibility (gym monitor)')
    p.add_argument('--monitor_path', default='/tmp/gym')
    p.add_argument('--disable_video', action='store_true')
    # Network
    p.add_argument('--network', default='simple', choices=['simple', 'cnn', 'nature'], help='Network architecture type')
    p.add_argument('--lr', default=0.001, type=float, help='Learning rate')
    p.add_argument('--reg_param', default=0.001, type=float, help='Regularization param')
    p.add_argument('--optimizer', default='sgd',
                   choices=['adadelta', 'adagrad', 'adam', 'ftrl', 'sgd', 'momentum', 'rmsprop'])
    p.add_argument('--momentum', default=0.9, type=float)
    p.add_argument('--rmsprop_decay', default=0.95, type=float)
    p.add_argument('--loss', default='mse', choices=['mse', 'huber'], help='mse = the reference; huber = north_star variant')
    # Agent
    p.add_argument('--num_episodes', default=10000, type=int)
    p.add_argument('--max_steps_per_episode', default=1000, type=int)
    p.add_argument('--minibatch_size', default=30, type=int)
    p.add_argument('--frames_per_state', default=1, type=int)
    p.add_argument('--resize_width', default=0, type=int)
    p.add_argument('--resize_height', default=0, type=int)
    p.add_argument('--reward_discount', default=0.9, type=float)
    p.add_argument('--replay_memory_capacity', default=10000, type=int)
    p.add_argument('--replay_start_size', default=0, type=int)
    p.add_argument('--init_random_action_prob', default=0.9, type=float)
    p.add_argument('--min_random_action_prob', default=0.1, type=float)
    p.add_argument('--random_action_explore_steps', default=10000, type=int)
    p.add_argument('--update_freq', default=1, type=int)
    p.add_argument('--target_update_freq', default=10000, type=int)
    # Distribution (accepted; rank/world size come from torchrun's environment)
    p.add_argument('--ps_hosts', default='')
    p.add_argument('--worker_hosts', default='localhost:0')
    p.add_argument('--job', default='worker', choices=['ps', 'worker'])
    p.add_argument('--task_id', default=0, type=int)
    p.add_argument('--gpu_id', default=0, type=int)
    p.add_argument('--sync', action='store_true')
    p.add_argument('--disable_cpu_param_pinning', action='store_true')
    p.add_argument('--disable_target_replication', action='store_true')
    # Summary
    p.add_argument('--logdir', default='/tmp/train_logs')
    p.add_argument('--summary_freq', default=100, type=int)
    # Extras
    p.add_argument('--seed', default=None, type=int)
    p.add_argument('--no_cuda_graph', dest='cuda_graph', action='store_false')
    p.add_argument('--precision', default=None, choices=['fp32', 'bf16'],
                   help='conv-net arithmetic; default bf16 for cnn/nature, fp32 for simple')
    p.add_argument('--dp', default='peer', choices=['peer', 'nccl'],
                   help='gradient exchange with more than one GPU: fused over NVLink peer memory, or NCCL all-reduce')
    args = p.parse_args(argv)
    if args.network == 'simple':
        args.precision = 'fp32'
    elif args.precision is None:
        args.precision = 'bf16'
    return args


class _Supervisor(object):
    """should_stop()-shaped object (the reference polls tf.train.Supervisor once per episode, dqn_agent.py:68-70)."""

    def __init__(self):
        import signal
        self._stop = False
        signal.signal(signal.SIGTERM, lambda *a: setattr(self, '_stop', True))

    def should_stop(self):
        return self._stop


def run_worker(args):
    import torch
    import envs
    from dqn_agent import DQNAgent
    from network import Network, Session
    from replay_memory import ReplayMemory
    import data_parallel
    rank, local, world = data_parallel.world_info()
    if world == 1:
        local = args.gpu_id
    torch.cuda.set_device(local)
    data_parallel.init(device=torch.device('cuda', local))
    env = envs.make(args.env, seed=data_parallel.shard_seed(args.seed, rank))
    input_shape = DQNAgent.get_input_shape(env, args)
    if args.precision == 'bf16' and len(input_shape) != 3:
        args.precision = 'fp32'                       # vector observations: nothing for the tensor cores to do
    network = Network.create_network(config=args, input_shape=input_shape,
                                     num_actions=env.action_space.n, num_replicas=world)
    # chief initialises, the others copy (main.py:133-143); mean-of-gradients scaling; exchange set-up
    data_parallel.setup_network(network, args, mode=args.dp)
    replay_memory = ReplayMemory(args.replay_memory_capacity, trusted=True)
    agent = DQNAgent(env, network, Session(), replay_memory, args, enable_summary=(rank == 0))
    agent.train(args.num_episodes, args.max_steps_per_episode, _Supervisor())
    if world > 1:
        import torch.distributed as dist
        agent._graphs.clear()                         # graphs that hold NCCL kernels go before the communicator
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    args = parse_args()
    logging.basicConfig(level=args.log_level, format='%(asctime)s %(levelname)s %(message)s')
    if args.job == 'ps':
        logging.warning('parameter servers are not used by the B200 path (NCCL data parallelism); exiting')
        sys.exit(0)
    run_worker(args)

#!/bin/bash
# Single-host multi-GPU DQN (the reference's scripts/dqn_multi_gpu.sh:1-105: same command line).
#   scripts/dqn_multi_gpu.sh <env_type> <env_name> <num_gpus>
# The reference starts one parameter server plus num_gpus-1 workers (GPU 0 only holds the parameters).  Here there
# is no parameter server: torchrun starts one worker per GPU, ALL num_gpus of them train, and the gradients are
# averaged synchronously (main.py --dp peer|nccl).  num_gpus <= 1 runs the single-GPU path.
if [ "$#" -ne 3 ]; then
  echo "Usage: $0 <env_type> <env_name> <num_gpus>.  Options for env_type: [control, atari]."
  exit 1
fi
HERE=$( cd "$(dirname "${BASH_SOURCE[0]}")" && pwd -P )
source "$HERE/dqn_params.sh"
params=$(dqn_params_for_env "$1" "$2") || exit 1
num_gpus=$3
LOG_ROOT="${DQN_LOG_ROOT:-/tmp}"
common="$params --logdir=$LOG_ROOT/train --monitor --monitor_path=$LOG_ROOT/gym --disable_video --sync $DQN_EXTRA_FLAGS"
if [[ "$num_gpus" -le 1 ]]; then
  echo "Starting 1 worker. Train logs: $LOG_ROOT/train"
  exec python "$HERE/../dist-dqn_b200/main.py" $common
fi
echo "Starting $num_gpus workers (one per GPU). Train logs: $LOG_ROOT/train"
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node="$num_gpus" --master-addr 127.0.0.1 \
  --master-port "${DQN_MASTER_PORT:-29533}" -- "$HERE/../dist-dqn_b200/main.py" $common      # "--": main.py's --monitor is not torchrun's

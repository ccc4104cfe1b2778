#!/bin/bash
# DQN on one B200 (the reference's scripts/dqn_single_node.sh:1-25: same command line, same presets).
#   scripts/dqn_single_node.sh <env_type> <env_name>        env_type: control | atari
# Extra main.py flags can be appended through DQN_EXTRA_FLAGS (e.g. "--precision fp32 --num_episodes 5").
if [ "$#" -ne 2 ]; then
  echo "Usage: $0 <env_type> <env_name>.  Options for env_type: [control, atari]."
  exit 1
fi
HERE=$( cd "$(dirname "${BASH_SOURCE[0]}")" && pwd -P )
source "$HERE/dqn_params.sh"
params=$(dqn_params_for_env "$1" "$2") || exit 1
LOG_ROOT="${DQN_LOG_ROOT:-/tmp}"
echo "Starting DQN on one GPU. Train logs: $LOG_ROOT/train"
exec python "$HERE/../dist-dqn_b200/main.py" $params --logdir="$LOG_ROOT/train" --monitor \
  --monitor_path="$LOG_ROOT/gym" --disable_video $DQN_EXTRA_FLAGS

#!/bin/bash
# Hyper-parameter presets of the two environment families, with the values the reference's
# scripts/dqn_params.sh:7-42 passes (they are the interface: bench.py and the tests use the same numbers).
# Usage (sourced): dqn_params_for_env <control|atari> <env_name>  -> prints the main.py flags.

declare -A DQN_PRESET_CONTROL=(
  [network]=simple [optimizer]=adam [lr]=0.001 [minibatch_size]=100 [num_episodes]=10000
  [max_steps_per_episode]=200 [replay_memory_capacity]=50000 [target_update_freq]=3000
  [reward_discount]=0.9 [init_random_action_prob]=0.5 [min_random_action_prob]=0.1
  [random_action_explore_steps]=50000
)
declare -A DQN_PRESET_ATARI=(
  [network]=cnn [optimizer]=rmsprop [lr]=0.00025 [minibatch_size]=32 [num_episodes]=10000
  [max_steps_per_episode]=500000 [replay_memory_capacity]=1000000 [target_update_freq]=10000
  [reward_discount]=0.99 [init_random_action_prob]=1.0 [min_random_action_prob]=0.1
  [random_action_explore_steps]=1000000 [frames_per_state]=4 [update_freq]=4
  [replay_start_size]=10000 [resize_width]=84 [resize_height]=84
)

dqn_params_for_env() {
  if [ "$#" -ne 2 ]; then
    echo "Usage: dqn_params_for_env <env_type> <env_name>; env_type is control or atari." >&2
    return 1
  fi
  local -n preset
  case "$1" in
    control) preset=DQN_PRESET_CONTROL ;;
    atari) preset=DQN_PRESET_ATARI ;;
    *) echo "Invalid env_type $1. Choices are [control, atari]." >&2; return 1 ;;
  esac
  local out="--env=$2"
  local key
  for key in "${!preset[@]}"; do out="$out --$key=${preset[$key]}"; done
  echo "$out"
}

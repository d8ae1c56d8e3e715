#!/bin/bash
# Developer tool: build variants of the actor kernel into tools/_lab/ (each a full libsalp_b200.so), then restore the default.
# usage: tools/build_actor_variants.sh name1:"-DX=1 -DY" name2:""   (the text after the colon becomes SALP_ACTOR_DEFS)
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/_lab
for spec in "$@"; do
  name="${spec%%:*}"; envs="${spec#*:}"
  touch underwater_swimmer_rl_b200/csrc/salp_actor.cu
  SALP_ACTOR_DEFS="$envs" python -m underwater_swimmer_rl_b200.build > /dev/null
  cp underwater_swimmer_rl_b200/libsalp_b200.so "tools/_lab/libsalp_${name}.so"
  echo "built $name ($envs)"
done
touch underwater_swimmer_rl_b200/csrc/salp_actor.cu
python -m underwater_swimmer_rl_b200.build > /dev/null

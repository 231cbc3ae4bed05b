#!/bin/sh
# step-kernel geometry sweep (development tool): envs per warp x CTA size, us per launch from tools/time_env.py
sizes=${1:-4096,65536,262144}
for epw in 4 8 16 32; do for blk in 32 64 128 256; do
  echo "== PF_EPW=$epw PF_BLOCK=$blk"
  PF_EPW=$epw PF_BLOCK=$blk python tools/time_env.py $sizes 2>&1 | grep post_physics
done; done

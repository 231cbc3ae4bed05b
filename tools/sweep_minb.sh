#!/bin/sh
# register-allocation variants of the step kernel (development tool)
for mb in 2 3 4; do for epw in 16 32; do
  echo "== PF_MINB=$mb PF_EPW=$epw"
  PF_MINB=$mb PF_EPW=$epw python tools/time_env.py ${1:-65536,262144} 2>&1 | grep post_physics
done; done

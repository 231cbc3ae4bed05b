#!/bin/bash
# A/B of the early weight staging of the inference kernels (PF_RL_EARLY_WEIGHTS=0 switches it off) on the bench step
line() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('%.5f ms/step  %.3e env-steps/s  act %.1f us' % (d['ms_per_step'], d['value'], d['roofline_act']['us_per_launch']))"; }
for h in 0 1 0 1; do echo "flat 4096  PF_RL_EARLY_WEIGHTS=$h"; PF_RL_EARLY_WEIGHTS=$h timeout 300 python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>/dev/null | line; done
for h in 0 1; do echo "rough 4096 PF_RL_EARLY_WEIGHTS=$h"; PF_RL_EARLY_WEIGHTS=$h timeout 300 python bench.py --steps 200 --warmup 10 --task pointfoot_rough --no-cpu-baseline 2>/dev/null | line; done

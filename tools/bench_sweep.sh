#!/bin/sh
# compact env-kernel sweep: task num_envs -> ms/step (graph), us per post_physics launch, roofline frac
for spec in "$@"; do
  task=${spec%%:*}; n=${spec##*:}
  python bench.py --steps 50 --warmup 5 --task $task --num-envs $n --no-cpu-baseline --no-policy | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$task', d['config']['num_envs'], 'ms/step=%.4f'%d['ms_per_step'], 'post_us=%.1f'%r['us_per_launch'], 'frac=%.3f'%r['frac'], 'value=%.3g'%d['value'], 'e2e=%.3g'%d['e2e']['value'])"
done

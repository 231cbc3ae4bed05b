#!/bin/bash
# round-end measurement pass on one B200 (run under gpurun): tests, bench lines, sweep, ncu launch list + full captures
O=gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python bench.py > $O/r2_bench_flat4096.json 2> $O/r2_bench.err; tail -2 $O/r2_bench.err
python bench.py --impl reference > $O/r2_bench_reference.json 2>> $O/r2_bench.err
: > $O/r2_bench_sweep.txt
for spec in pointfoot_flat:4096 pointfoot_flat:16384 pointfoot_flat:65536 pointfoot_flat:262144 pointfoot_rough:4096 pointfoot_rough:65536 pointfoot_rough:262144; do
  task=${spec%%:*}; n=${spec##*:}
  python bench.py --steps 100 --warmup 5 --task $task --num-envs $n --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('%-16s %7d  ms/step=%.4f value=%.3e e2e=%.3e post_us=%.1f frac=%.3f' % ('$task', d['config']['num_envs'], d['ms_per_step'], d['value'], d['e2e']['value'], r['us_per_launch'], r['frac']))" | tee -a $O/r2_bench_sweep.txt
done
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/b.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_bench_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pf_post_physics -s 20 -c 1 -f -o $O/r2_post_flat4096 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/ncu_a.log 2>&1
python tools/time_env.py 262144 > $O/time_env_262k.txt 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:pf_post_physics|pf_finalize" -s 18 -c 2 -f -o $O/r2_post_flat262k python tools/time_env.py 262144 pointfoot_flat > $O/ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:pf_post_physics|pf_finalize|pf_height_scan" -s 27 -c 3 -f -o $O/r2_post_rough262k python tools/time_env.py 262144 pointfoot_rough > $O/ncu_c.log 2>&1
cat $O/time_env_262k.txt | grep post

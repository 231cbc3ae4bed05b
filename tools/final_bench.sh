O=gpurun_out
python bench.py > $O/r2_bench_flat4096.json 2> $O/r2_bench.err; tail -1 $O/r2_bench.err
python bench.py --impl reference > $O/r2_bench_reference.json 2>> $O/r2_bench.err
python bench.py --task pointfoot_rough > $O/r2_bench_rough4096.json 2>> $O/r2_bench.err
: > $O/r2_bench_sweep.txt
for spec in pointfoot_flat:4096 pointfoot_flat:65536 pointfoot_flat:262144 pointfoot_rough:4096 pointfoot_rough:65536 pointfoot_rough:262144; do
  task=${spec%%:*}; n=${spec##*:}
  python bench.py --steps 100 --warmup 5 --task $task --num-envs $n --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('%-16s %7d  ms/step=%.4f value=%.3e e2e=%.3e post_us=%.1f frac=%.3f act_us=%s' % ('$task', d['config']['num_envs'], d['ms_per_step'], d['value'], d['e2e']['value'], r['us_per_launch'], r['frac'], d.get('policy',{}).get('us_per_launch')))" | tee -a $O/r2_bench_sweep.txt
done

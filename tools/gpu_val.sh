# device-timed lines only (value of every workload), for kernel A/B
export PYTHONPATH=.
mkdir -p gpurun_out
for pf in 1; do timeout 900 python bench.py --no-p3p --no-cpu --no-e2e --no-cpu-case --steps 50 > gpurun_out/val.json 2> gpurun_out/val.err; echo "rc $?"
grep '^{' gpurun_out/val.json | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('cfg4', d['ms_per_step'], d['roofline']['frac'], 'ok', d['all_frames_accepted'], 'tracking', d['tracking']['ms_per_step'])
for k in ('single_filter','outlier_heavy','sharded_filter'):
    s=d.get(k) or {}
    print('  ', k, 'ms/step', s.get('ms_per_step'), 'frac', (s.get('roofline') or {}).get('frac'))
"
done

# all-gather (LL) grid reduction: parity subset, then A/B of the latency-bound lines against the atomic path (PFMPE_LL=0)
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_cuda_parity.py tests/test_cuda_track.py tests/test_cuda_shard.py -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/r02q_pytest.log 2>&1; tail -5 gpurun_out/r02q_pytest.log
for ll in 1 0; do
  PFMPE_LL=$ll timeout 900 python bench.py --no-p3p --no-cpu --steps 20 --no-e2e > gpurun_out/r02q_ll$ll.json 2> gpurun_out/r02q_ll$ll.err; echo "LL=$ll rc $?"
  grep '^{' gpurun_out/r02q_ll$ll.json | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('single_filter','outlier_heavy','cpu_case','sharded_filter'):
    s=d.get(k) or {}
    print('  ', k, 'ms/step', s.get('ms_per_step'), 'p50', s.get('latency_p50_ms'), 'frac', (s.get('roofline') or {}).get('frac'))
"
done

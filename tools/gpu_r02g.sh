set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-r02g}
timeout 1200 python -m pytest tests -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/${TAG}_pytest.log 2>&1; tail -25 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -5 gpurun_out/${TAG}_bench.err; python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
def brief(o):
    return {k:(v if not isinstance(v,dict) else {kk:vv for kk,vv in v.items() if kk in ("value","frac","ms_per_step","p50_ms","achieved")}) for k,v in o.items() if k in ("value","ms_per_step","latency_p50_ms","roofline","e2e","cpu_baseline","tracking","state_checksums","device_ms_per_step")}
print("HEAD", brief(d))
for k in ("single_filter","outlier_heavy","cpu_case","sharded_filter","p3p_vote"):
    if k in d: print(k, brief(d[k]))
PY

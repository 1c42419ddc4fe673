set -x
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_track.py -m gpu -x -q > gpurun_out/r02n_pytest.log 2>&1; tail -5 gpurun_out/r02n_pytest.log
true
python - <<PY
import json
d=json.load(open("gpurun_out/r02n_bench.json"))
print("value", d["value"], d["ms_per_step"], "frac", d["roofline"]["frac"]); print("tracking", d["tracking"]); print("e2e", {k:v for k,v in d["e2e"].items() if k!="path"})
PY

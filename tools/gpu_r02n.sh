set -x
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_track.py -m gpu -x -q > gpurun_out/r02n_pytest.log 2>&1; tail -5 gpurun_out/r02n_pytest.log
PFMPE_TRACE=1 timeout 900 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 30 > gpurun_out/r02n_bench.json 2> gpurun_out/r02n_bench.err; grep "pfmpe trace" gpurun_out/r02n_bench.err | tail -2
python - <<PY
import json
d=json.load(open("gpurun_out/r02n_bench.json"))
print("value", d["value"], d["ms_per_step"], "frac", d["roofline"]["frac"]); print("tracking", d["tracking"]); print("e2e", {k:v for k,v in d["e2e"].items() if k!="path"})
PY

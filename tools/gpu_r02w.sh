export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 50 > gpurun_out/r02w.json 2> gpurun_out/r02w.err; echo "rc $?"; tail -3 gpurun_out/r02w.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02w.json"))
print("value", d["value"], d["ms_per_step"], "frac", d["roofline"]["frac"]); print("single_launch", d["single_launch"]); print("tracking", d["tracking"]); print("e2e", {k:v for k,v in d["e2e"].items() if k!="path"})
PY

# the device-tail batch call: its tests, then the bench e2e line with the per-phase trace
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_cuda_track.py -m gpu -x -q > gpurun_out/e2e_pytest.log 2>&1; tail -4 gpurun_out/e2e_pytest.log
for ov in 1; do
PFMPE_TRACE=1 timeout 900 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 30 > gpurun_out/e2e_bench.json 2> gpurun_out/e2e_bench.err; echo "rc $?"
grep "pfmpe trace" gpurun_out/e2e_bench.err | tail -1 | cut -c1-400
python - <<PY
import json
d=json.load(open("gpurun_out/e2e_bench.json"))
print("   e2e", d["e2e"]["ms_per_step"], d["e2e"]["p50_ms"], d["e2e"]["value"], "host_tail", d["e2e"]["host_tail"]["ms_per_step"], "ok", d["all_frames_accepted"])
PY
done

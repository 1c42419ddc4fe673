export PYTHONPATH=.
mkdir -p gpurun_out
true
for sp in 1; do
PFMPE_SPLIT_TAIL=$sp PFMPE_TRACE=1 timeout 900 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 30 > gpurun_out/r02x_$sp.json 2> gpurun_out/r02x_$sp.err; echo "split=$sp rc $?"
grep "pfmpe trace" gpurun_out/r02x_$sp.err | tail -1
python - <<PY
import json
d=json.load(open("gpurun_out/r02x_$sp.json"))
print("   value", d["ms_per_step"], "tracking", d["tracking"]["ms_per_step"], d["tracking"]["single_launch"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["p50_ms"], "ok", d["all_frames_accepted"])
PY
done

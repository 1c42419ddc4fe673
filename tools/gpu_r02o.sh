# device-tail batch e2e against the Gauss-Newton cap (PFMPE_GN_CAP): trace medians + the bench e2e line
export PYTHONPATH=.
mkdir -p gpurun_out
nproc
for cap in 16 32 64 128; do
  PFMPE_GN_CAP=$cap PFMPE_TRACE=1 timeout 900 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 30 > gpurun_out/r02o_$cap.json 2> gpurun_out/r02o_$cap.err
  echo "cap $cap"; grep "pfmpe trace" gpurun_out/r02o_$cap.err | tail -1
  python - <<PY
import json
d=json.load(open("gpurun_out/r02o_$cap.json"))
e=d["e2e"]; print("   e2e ms/step", e["ms_per_step"], "p50", e["p50_ms"], "PME/s", e["value"])
PY
done

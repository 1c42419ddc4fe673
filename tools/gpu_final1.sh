# final numbers on one B200: the default bench (as the driver runs it) and the reference arm
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1500 python bench.py > gpurun_out/final1_bench.raw 2> gpurun_out/final1_bench.err; echo "bench rc $?"
grep '^{' gpurun_out/final1_bench.raw | tail -1 > gpurun_out/final1_bench.json; tail -2 gpurun_out/final1_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/final1_ref.raw 2> gpurun_out/final1_ref.err; echo "reference rc $?"
grep '^{' gpurun_out/final1_ref.raw | tail -1 > gpurun_out/final1_ref.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/final1_bench.json"))
print("HEAD value %.4e ms %.4f frac %.3f launches %s clocks %s" % (d["value"], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"], d.get("clocks")))
print("tracking", d["tracking"]["ms_per_step"], d["tracking"]["single_launch"]["ms_per_step"], "single_launch", d["single_launch"]["ms_per_step"])
print("e2e", {k: v for k, v in d["e2e"].items() if k not in ("path", "host_tail")}, "host_tail", d["e2e"].get("host_tail", {}).get("ms_per_step"))
print("cpu_baseline", d.get("cpu_baseline"))
for k in ("single_filter", "outlier_heavy", "cpu_case", "sharded_filter", "p3p_vote"):
    if k in d:
        s = d[k]; print(k, "%.4f ms" % s["ms_per_step"], "%.3e" % s["value"], (s.get("roofline") or {}).get("frac"), {kk: vv for kk, vv in (s.get("e2e") or {}).items() if kk in ("value", "ms_per_step", "p50_ms", "ms")}, (s.get("cpu_baseline") or {}).get("value"))
r = json.load(open("gpurun_out/final1_ref.json"))
print("REF", {k: r.get(k) for k in ("impl", "value", "unit", "ms_per_step", "steps")}, r.get("cpu_baseline"))
print("e2e ratio vs reference arm: %.1f" % (d["e2e"]["value"] / r["value"]))
PY

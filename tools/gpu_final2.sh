# final-state check on 2 GPUs: the whole GPU suite (2-process test included), smoke(), bench at N=2 exactly as the driver launches it
export PYTHONPATH=.
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/final2_pytest.log 2>&1; tail -4 gpurun_out/final2_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final2_smoke.log 2>&1; echo "smoke rc $?"; tail -3 gpurun_out/final2_smoke.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/final2_bench.raw 2> gpurun_out/final2_bench.err; echo "bench rc $?"
grep '^{' gpurun_out/final2_bench.raw | tail -1 > gpurun_out/final2_bench.json; tail -3 gpurun_out/final2_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/final2_bench.json"))
print("HEAD value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "launches", d["gpu_launches"], "wall", d.get("wall_s"))
print("e2e", {k: v for k, v in d["e2e"].items() if k != "path"})
for k in ("single_filter", "outlier_heavy", "cpu_case", "sharded_filter", "p3p_vote"):
    if k in d:
        s = d[k]; print(k, s.get("ms_per_step"), s.get("value"), (s.get("roofline") or {}).get("frac"), {kk: vv for kk, vv in (s.get("e2e") or {}).items() if kk in ("value", "ms_per_step", "p50_ms")}, s.get("bit_identical"))
PY

# bench at N GPUs exactly as the driver launches it
export PYTHONPATH=.
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/final${N}_bench.raw 2> gpurun_out/final${N}_bench.err; echo "bench rc $?"
grep '^{' gpurun_out/final${N}_bench.raw | tail -1 > gpurun_out/final${N}_bench.json; tail -3 gpurun_out/final${N}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/final${N}_bench.json"))
print("HEAD value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "launches", d["gpu_launches"])
print("e2e", {k: v for k, v in d["e2e"].items() if k != "path"})
for k in ("single_filter", "outlier_heavy", "cpu_case", "sharded_filter", "p3p_vote"):
    if k in d:
        s = d[k]; print(k, s.get("ms_per_step"), s.get("value"), (s.get("roofline") or {}).get("frac"), s.get("bit_identical"), s.get("state_checksums"))
PY

# the sharded sub-line alone at N ranks: its whole-filter checksums must not depend on N
export PYTHONPATH=.
mkdir -p gpurun_out
N=${1:-2}
F="--no-single --no-outlier --no-cpu-case --no-p3p --no-cpu --no-e2e --steps 5 --warmup 3"
timeout 600 python bench.py $F > gpurun_out/cks1.raw 2> gpurun_out/cks1.err; echo "N=1 rc $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N $F > gpurun_out/cksN.raw 2> gpurun_out/cksN.err; echo "N=$N rc $?"
python - <<'PY'
import json
for f in ("gpurun_out/cks1.raw", "gpurun_out/cksN.raw"):
    line = [l for l in open(f) if l.startswith("{")][-1]
    d = json.loads(line); s = d["sharded_filter"]
    print(d["n_gpus"], s["ms_per_step"], s["state_checksums"])
PY

# multi-GPU checks: 2-process sharded bit-identity test, sharded filter timing (C loop / per-stage python loop), bench
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
N=${1:-2}; TAG=${2:-r02_n2}
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_cuda_shard_mp.py -m gpu -x -q > gpurun_out/${TAG}_pytest_mp.log 2>&1; tail -5 gpurun_out/${TAG}_pytest_mp.log
timeout 300 $TR --master-port 29511 tools/shard_nccl_check.py --particles 262144 --frames 8 2>&1 | grep shard_nccl | tee gpurun_out/${TAG}_shard_check.json
timeout 300 $TR --master-port 29512 tools/shard_nccl_check.py --particles 67108864 --frames 14 --no-compare 2>&1 | grep shard_nccl | tee -a gpurun_out/${TAG}_shard_check.json
timeout 300 $TR --master-port 29513 tools/shard_nccl_check.py --particles 67108864 --frames 10 --no-compare --stages 2>&1 | grep "frame\|shard_nccl" | cut -c1-330 | tee gpurun_out/${TAG}_shard_stages.txt
timeout 900 $TR --master-port 29514 bench.py --gpus $N --no-outlier --no-cpu-case --no-p3p > gpurun_out/${TAG}_bench.raw 2> gpurun_out/${TAG}_bench.err; grep '^{' gpurun_out/${TAG}_bench.raw | tail -1 > gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
def brief(o):
    return {k:(v if not isinstance(v,dict) else {kk:vv for kk,vv in v.items() if kk in ("value","frac","ms_per_step","p50_ms","achieved")}) for k,v in o.items() if k in ("value","ms_per_step","latency_p50_ms","roofline","e2e","tracking","state_checksums","device_ms_per_step","n_gpus")}
print("HEAD", brief(d))
for k in ("single_filter","sharded_filter"):
    if k in d: print(k, brief(d[k]))
PY

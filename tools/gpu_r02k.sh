set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-r02k}
timeout 1200 python -m pytest tests -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/${TAG}_pytest.log 2>&1; tail -25 gpurun_out/${TAG}_pytest.log
timeout 300 python tools/shard_nccl_check.py --particles 67108864 --frames 10 --no-compare 2>&1 | grep shard_nccl
timeout 300 python tools/shard_nccl_check.py --particles 67108864 --frames 8 --no-compare --stages 2>&1 | grep "frame\|shard_nccl"

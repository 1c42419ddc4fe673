# sharded filter only, N ranks: bit-identity of the block-cyclic placement, then timing at 2^26 through the C loop
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
N=${1:-8}; TAG=${2:-r02_n8s}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29520
for v in 4 2; do P=$((P+1)); timeout 300 $TR --master-port $P tools/shard_nccl_check.py --particles 4194304 --frames 10 --shards-per-gpu $v 2>&1 | grep shard_nccl | cut -c1-260 | tee -a gpurun_out/${TAG}_shard_check.json; done
for v in 1 2 4; do P=$((P+1)); timeout 300 $TR --master-port $P tools/shard_nccl_check.py --particles 67108864 --frames 40 --no-compare --shards-per-gpu $v 2>&1 | grep shard_nccl | tee -a gpurun_out/${TAG}_shard_check.json; done
P=$((P+1)); PFMPE_SCATTER=sparse timeout 300 $TR --master-port $P tools/shard_nccl_check.py --particles 67108864 --frames 40 --no-compare --shards-per-gpu 1 2>&1 | grep shard_nccl | sed 's/^/SPARSE /' | tee -a gpurun_out/${TAG}_shard_check.json

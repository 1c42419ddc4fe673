export PYTHONPATH=.
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/shard_nccl_check.py --particles 1048576 --frames 6 --shards-per-gpu 4 > gpurun_out/dbg.log 2>&1; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/dbg.log | tail -25

set -x
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cuda_shard.py tests/test_cuda_fullsize.py -m gpu -x -q -k "shard or config5 or checksum" > gpurun_out/shardtest.log 2>&1; tail -6 gpurun_out/shardtest.log

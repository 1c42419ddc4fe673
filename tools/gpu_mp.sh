# the 2-process / 2-GPU tests of the sharded filter alone
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_cuda_shard_mp.py -m gpu -x -q > gpurun_out/mp_pytest.log 2>&1; tail -4 gpurun_out/mp_pytest.log

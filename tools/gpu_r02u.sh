export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_cuda_shard.py -m gpu -x -q > gpurun_out/r02u_pytest.log 2>&1; tail -15 gpurun_out/r02u_pytest.log

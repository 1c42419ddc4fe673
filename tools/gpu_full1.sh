# whole GPU suite + smoke + the default bench on one B200 (the 2-process test skips)
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/full1_pytest.log 2>&1; tail -4 gpurun_out/full1_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/full1_smoke.log 2>&1; echo "smoke rc $?"; tail -3 gpurun_out/full1_smoke.log
bash tools/gpu_final1.sh

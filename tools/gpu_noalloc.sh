set -x
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; tail -4 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_suite.log 2>&1; tail -6 gpurun_out/noalloc_suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4

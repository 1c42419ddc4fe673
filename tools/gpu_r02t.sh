export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cuda_p3p.py tests/test_zz_no_alloc.py tests/test_pose_estimator_init.py -m gpu -x -q > gpurun_out/r02t_pytest.log 2>&1; tail -4 gpurun_out/r02t_pytest.log
python tools/p3p_one.py 205 5; python tools/p3p_one.py 105 5; python tools/p3p_one.py 45 5
echo "single kernel:"; PFMPE_P3P_ROOTS_MB=0 python tools/p3p_one.py 205 5; PFMPE_P3P_ROOTS_MB=0 python tools/p3p_one.py 105 5
echo "roots occ 6 / vote occ 8:"; PFMPE_P3P_OCC_ROOTS=6 python tools/p3p_one.py 205 5
echo "roots occ 8 / vote occ 6:"; PFMPE_P3P_OCC_ROOTS=8 PFMPE_P3P_OCC=6 python tools/p3p_one.py 205 5
echo "slabs of 512 MB:"; PFMPE_P3P_ROOTS_MB=512 python tools/p3p_one.py 205 5

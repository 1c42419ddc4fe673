export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cuda_p3p.py tests/test_zz_no_alloc.py tests/test_pose_estimator_init.py -m gpu -x -q > gpurun_out/r02t_pytest.log 2>&1; tail -4 gpurun_out/r02t_pytest.log
for occ in 5 6 8; do echo "vote kernel occ $occ (roots kernel 8):"; PFMPE_P3P_OCC_ROOTS=8 PFMPE_P3P_OCC=$occ python tools/p3p_one.py 205 5; done
python tools/p3p_one.py 45 5
echo "single kernel"; PFMPE_P3P_ROOTS_MB=0 python tools/p3p_one.py 205 5

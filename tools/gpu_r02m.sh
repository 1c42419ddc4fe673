set -x
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err; tail -3 gpurun_out/r02m_bench.err
bash tools/gpu_profile_r02.sh > gpurun_out/r02m_profile.log 2>&1; tail -12 gpurun_out/r02m_profile.log

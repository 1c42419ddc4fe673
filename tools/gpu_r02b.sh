# round 2: after the instruction diet of the score / resample stages -- tests, timeline, source-level captures
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-r02b}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -15 gpurun_out/${TAG}_pytest.log
timeout 300 python tools/timeline.py > gpurun_out/${TAG}_timeline.txt 2>&1; cat gpurun_out/${TAG}_timeline.txt
timeout 300 python tools/one_cfg.py 4096 1024 10 6 > gpurun_out/one_cfg4.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 3 -c 1 -o gpurun_out/${TAG}_full_cfg4 \
    python tools/one_cfg.py 4096 1024 10 6 > gpurun_out/ncu_cfg4.log 2>&1
timeout 300 python tools/one_cfg.py 16777216 1 10 6 > gpurun_out/one_cfg5s.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 3 -c 1 -o gpurun_out/${TAG}_full_cfg5s \
    python tools/one_cfg.py 16777216 1 10 6 > gpurun_out/ncu_cfg5s.log 2>&1
cat gpurun_out/one_cfg4.log gpurun_out/one_cfg5s.log

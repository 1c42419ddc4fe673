# round 2: replay-mode measurements (what bench.py times) + source-level captures, dev build (L = 5 only)
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-r02c}
timeout 900 python -m pytest tests -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/${TAG}_pytest.log 2>&1; tail -15 gpurun_out/${TAG}_pytest.log
for a in "4096 4096 10 12" "100000 1 10 40 1" "1000000 1 200 12 1" "16777216 1 10 8" "67108864 1 10 6"; do timeout 300 python tools/replay_cfg.py $a; done 2>&1 | tee gpurun_out/${TAG}_replay.txt
timeout 300 python tools/replay_cfg.py 4096 1024 10 6 > gpurun_out/one_cfg4.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 9 -c 1 -o gpurun_out/${TAG}_full_cfg4 \
    python tools/replay_cfg.py 4096 1024 10 6 > gpurun_out/ncu_cfg4.log 2>&1
timeout 300 python tools/replay_cfg.py 16777216 1 10 6 > gpurun_out/one_cfg5s.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 9 -c 1 -o gpurun_out/${TAG}_full_cfg5s \
    python tools/replay_cfg.py 16777216 1 10 6 > gpurun_out/ncu_cfg5s.log 2>&1
tail -3 gpurun_out/ncu_cfg4.log

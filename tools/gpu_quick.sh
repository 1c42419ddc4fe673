set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-q}
timeout 900 python -m pytest tests -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/${TAG}_pytest.log 2>&1; tail -5 gpurun_out/${TAG}_pytest.log
for a in "4096 4096 10 12" "100000 1 10 40 1" "1000000 1 200 12 1" "16777216 1 10 8" "67108864 1 10 6"; do timeout 300 python tools/replay_cfg.py $a; done 2>&1 | tee gpurun_out/${TAG}_replay.txt
timeout 300 python tools/timeline.py > gpurun_out/${TAG}_timeline.txt 2>&1; cat gpurun_out/${TAG}_timeline.txt

set -x
export PYTHONPATH=.
mkdir -p gpurun_out
TAG=${1:-r02j}
timeout 1200 python -m pytest tests -m gpu -x -q -k "not edge_cases and not smallest and not leds" > gpurun_out/${TAG}_pytest.log 2>&1; tail -25 gpurun_out/${TAG}_pytest.log
for c in 0 1; do for t in 512 1024 4096; do PFMPE_CLUSTER=$c timeout 300 python tools/replay_cfg.py 4096 $t 10 12 | cut -c1-60,200-400; done; done 2>&1 | tee gpurun_out/${TAG}_cluster_ab.txt

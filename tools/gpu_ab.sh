export PYTHONPATH=.
mkdir -p gpurun_out
for c in 0 1; do for t in 512 1024 4096; do PFMPE_CLUSTER=$c timeout 300 python tools/replay_cfg.py 4096 $t 10 12 | sed 's/ms\/frame:.*| p50/| p50/'; done; done 2>&1 | tee gpurun_out/r02j_cluster_ab.txt

# round 2 profile recipe (B200_PROFILING.md): launch list of the bench command, full captures of the headline kernel
# (cfg4 batch shape), the grid shape (cfg2, cfg3, 2^24) and the stage-only launches of the sharded filter (score / resample).
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
B="--steps 10 --warmup 3 --only --no-cpu --no-e2e"
python bench.py $B > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_cfg4.csv python bench.py $B > gpurun_out/ncu_launches.log 2>&1
python tools/replay_cfg.py 4096 4096 10 6 > gpurun_out/one_cfg4.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 9 -c 1 -o gpurun_out/r02_full_cfg4 python tools/replay_cfg.py 4096 4096 10 6 > gpurun_out/ncu_cfg4.log 2>&1
python tools/replay_cfg.py 100000 1 10 12 1 > gpurun_out/one_cfg2.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 18 -c 1 -o gpurun_out/r02_full_cfg2 python tools/replay_cfg.py 100000 1 10 12 1 > gpurun_out/ncu_cfg2.log 2>&1
python tools/replay_cfg.py 1000000 1 200 8 1 > gpurun_out/one_cfg3.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 12 -c 1 -o gpurun_out/r02_full_cfg3 python tools/replay_cfg.py 1000000 1 200 8 1 > gpurun_out/ncu_cfg3.log 2>&1
python tools/replay_cfg.py 16777216 1 10 6 > gpurun_out/one_cfg5s.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 9 -c 1 -o gpurun_out/r02_full_cfg5s python tools/replay_cfg.py 16777216 1 10 6 > gpurun_out/ncu_cfg5s.log 2>&1
python tools/shard_stage_one.py 16777216 5 > gpurun_out/one_stages.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 6 -c 3 -o gpurun_out/r02_full_stages python tools/shard_stage_one.py 16777216 5 > gpurun_out/ncu_stages.log 2>&1
cat gpurun_out/one_*.log | cut -c1-60,300-500; tail -2 gpurun_out/ncu_stages.log

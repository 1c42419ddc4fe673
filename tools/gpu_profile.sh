# Profile recipe of B200_PROFILING.md for the bench workload (run under gpurun, one GPU).
set -x
export PYTHONPATH=.
mkdir -p gpurun_out
B="--steps 30 --warmup 5 --no-cpu --no-e2e --no-batched --no-sharded --no-p3p --no-outlier"
python bench.py $B > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg2.csv \
    python bench.py $B > gpurun_out/ncu_launches.log 2>&1
# full capture of one replay launch (the timed kernel): skip the tracked run (35 launches) + warm-up
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 50 -c 1 -o gpurun_out/full_cfg2 \
    python bench.py $B > gpurun_out/ncu_full_cfg2.log 2>&1
ncu -i gpurun_out/full_cfg2.ncu-rep --page raw --csv > gpurun_out/full_cfg2_raw.csv 2>/dev/null
python tools/one_cfg.py 16777216 > gpurun_out/one_cfg5s.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:pf_frames -s 4 -c 1 -o gpurun_out/full_cfg5s \
    python tools/one_cfg.py 16777216 > gpurun_out/ncu_full_cfg5s.log 2>&1
ncu -i gpurun_out/full_cfg5s.ncu-rep --page raw --csv > gpurun_out/full_cfg5s_raw.csv 2>/dev/null
# K9 (row f4): the initialisation vote, 205 detections
python tools/p3p_one.py 205 3 > gpurun_out/one_p3p.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_p3p.csv \
    python tools/p3p_one.py 205 3 > gpurun_out/ncu_launches_p3p.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:pf_p3p_hist -s 1 -c 1 -o gpurun_out/full_p3p \
    python tools/p3p_one.py 205 2 > gpurun_out/ncu_full_p3p.log 2>&1
ncu -i gpurun_out/full_p3p.ncu-rep --page raw --csv > gpurun_out/full_p3p_raw.csv 2>/dev/null
tail -2 gpurun_out/ncu_full_cfg2.log gpurun_out/ncu_full_cfg5s.log gpurun_out/ncu_full_p3p.log; cat gpurun_out/one_cfg5s.log gpurun_out/one_p3p.log

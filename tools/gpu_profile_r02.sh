# round 2 profile recipe (B200_PROFILING.md): launch list of the bench command, full captures of the headline kernel
# (cfg4 batch shape), the grid shape (cfg2, cfg3, 2^24) and the stage-only launches of the sharded filter (score / resample).
# Every report is reduced to text on the box (counter summary, raw page, per-stage instruction mix) and then deleted:
# gpurun brings back at most 64 MiB.
set -x
export PYTHONPATH=.
O=gpurun_out/prof
rm -rf $O; mkdir -p $O
B="--steps 10 --warmup 3 --only --no-cpu --no-e2e"
python bench.py $B > $O/prof_bench.json 2> $O/prof_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r02_launches_cfg4.csv python bench.py $B > $O/ncu_launches.log 2>&1
cap() {  # name, skip, count, steps-for-sass_hist, command...
  name=$1; skip=$2; cnt=$3; steps=$4; shift 4
  "$@" > $O/one_$name.log 2>&1 &&
  ncu --set full --import-source on --clock-control none -k regex:pf_frames -s $skip -c $cnt -o $O/r02_full_$name "$@" > $O/ncu_$name.log 2>&1
  python tools/profile_summary.py $O/r02_full_$name.ncu-rep > $O/r02_ncu_summary_$name.txt 2>&1
  ncu -i $O/r02_full_$name.ncu-rep --page raw --csv > $O/r02_ncu_full_${name}_raw.csv 2>/dev/null
  if [ "$steps" != "0" ]; then python tools/sass_hist.py $O/r02_full_$name.ncu-rep $steps 40 > $O/r02_sass_hist_$name.txt 2>&1; fi
  rm -f $O/r02_full_$name.ncu-rep
}
cap cfg4 9 1 524288 python tools/replay_cfg.py 4096 4096 10 6
cap cfg2 18 1 3125 python tools/replay_cfg.py 100000 1 10 12 1
cap cfg3 12 1 31250 python tools/replay_cfg.py 1000000 1 200 8 1
cap cfg5s 9 1 524288 python tools/replay_cfg.py 16777216 1 10 6
cap stages 6 3 0 python tools/shard_stage_one.py 16777216 5
cat $O/one_*.log | cut -c1-60,300-500; du -sh gpurun_out

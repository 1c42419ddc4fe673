# one full ncu capture of the K9 vote kernel at D = 205 (after the same command ran clean), reduced to text on the box
export PYTHONPATH=.
O=gpurun_out/prof_p3p; rm -rf $O; mkdir -p $O
python tools/p3p_one.py 205 3 > $O/one.log 2>&1 && cat $O/one.log &&
ncu --set full --import-source on --clock-control none -k regex:pf_p3p_hist -s 0 -c 2 -o $O/full python tools/p3p_one.py 205 2 > $O/ncu.log 2>&1
python tools/profile_summary.py $O/full.ncu-rep > $O/r02_ncu_summary_p3p.txt 2>&1
ncu -i $O/full.ncu-rep --page raw --csv > $O/r02_ncu_full_p3p_raw.csv 2>/dev/null
python tools/sass_hist.py $O/full.ncu-rep 2652956 50 > $O/r02_sass_hist_p3p.txt 2>&1
rm -f $O/full.ncu-rep
cat $O/r02_ncu_summary_p3p.txt

mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
python bench.py > gpurun_out/r1_bench_cfg2.json 2> gpurun_out/r1_bench_cfg2.err
python tools/timeline.py > gpurun_out/r1_timeline.txt 2>&1
tail -3 gpurun_out/r1_pytest.log; cat gpurun_out/r1_bench_cfg2.json; cat gpurun_out/r1_timeline.txt

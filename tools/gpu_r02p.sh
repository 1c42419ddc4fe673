# track tests, then the launch list of the bench command (frame / tail kernels only: the setup fills are filtered out)
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_track.py -m gpu -x -q > gpurun_out/r02p_pytest.log 2>&1; tail -5 gpurun_out/r02p_pytest.log
CMD="python bench.py --no-cpu --steps 10 --warmup 3"
timeout 900 $CMD > gpurun_out/r02p_bench.json 2> gpurun_out/r02p_bench.err; echo "bench rc $?"
grep '^{' gpurun_out/r02p_bench.json | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value', d['value'], d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], d['e2e'].get('ms_per_step'))
for s in d.get('sub', d.get('lines', [])) if isinstance(d.get('sub', d.get('lines', [])), list) else []:
    print('  ', s.get('workload'), s.get('ms_per_step'), s.get('roofline', {}).get('frac'), (s.get('e2e') or {}).get('ms_per_step'))
"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:pf_(frames|track|shard|p3p|meancov|roi|checksum)' -c 3000 --csv --log-file gpurun_out/r02p_launches.csv $CMD > gpurun_out/r02p_ncu.log 2>&1; echo "ncu rc $?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/r02p_launches.csv")) if len(r) > 10]
hdr = rows[0]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value"); u = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    t = float(r[v].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[u], 1.0)
    name = r[k].split("(")[0]
    a = agg.setdefault((name, r[hdr.index("Grid Size")], r[hdr.index("Block Size")]), [0, 0.0]); a[0] += 1; a[1] += t
tot = sum(a[1] for a in agg.values())
with open("gpurun_out/r02p_launch_summary.txt", "w") as f:
    for (n, g, b), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        line = "%-60s grid %-18s block %-14s launches %5d  total %10.1f us  mean %9.1f us  share %5.1f%%" % (n[:60], g, b, c, t, t / c, 100 * t / tot)
        print(line); f.write(line + "\n")
PY

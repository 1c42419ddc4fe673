# device-tail batch e2e against the device Gauss-Newton cap (split tail kernel, 3x3-block host form)
export PYTHONPATH=.
mkdir -p gpurun_out
for cap in 24 32 48 64 96; do
PFMPE_GN_CAP=$cap PFMPE_TRACE=1 timeout 600 python bench.py --no-single --no-outlier --no-cpu-case --no-p3p --no-sharded --no-cpu --steps 20 > gpurun_out/cap.json 2> gpurun_out/cap.err
echo "cap $cap: $(grep 'pfmpe trace' gpurun_out/cap.err | tail -1 | sed 's/.*host GN/host GN/' | cut -c1-200)"
python -c "
import json; d=json.load(open('gpurun_out/cap.json')); e=d['e2e']; print('    e2e ms/step %.4f p50 %.4f' % (e['ms_per_step'], e['p50_ms']))"
done

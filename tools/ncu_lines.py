"""Aggregate an ncu --page source --print-source cuda,sass CSV per source line (stall samples, instructions)."""
import csv, collections, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; cur = None; agg = collections.OrderedDict(); ts = ti = 0
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if r[0] in ('Function Name', 'Kernel Name'): continue
    if r[0] != '' and hdr:
        d = dict(zip(hdr, r))
        def num(x):
            try: return float(x)
            except: return 0.0
        s = num(r[4]); i = num(d.get('Instructions Executed', 0))
        a = agg.setdefault((cur, int(r[0]), r[1][:100]), [0, 0]); a[0] += s; a[1] += i; ts += s; ti += i
print("total samples %d, warp instructions %.4g" % (ts, ti))
for (f, ln, src), (s, i) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% ins  %s:%d  %s" % (100 * s / ts, 100 * i / ti, f, ln, src))

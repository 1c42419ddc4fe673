"""Static SASS size of pf_frames_kernel<5> by source function (nvdisasm line info; inlined line = innermost)."""
import sys, re, collections, subprocess, os, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "pf_monocular_pose_estimator_b200", "libpfmpe.so")], cwd=tmp, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
insec = False; cur = None; cnt = collections.Counter()
for l in out.splitlines():
    if l.startswith("//---") and ".text." in l:
        insec = "pf_frames_kernelILi5ELb0" in l or "pf_frame_tailILi5ELb0" in l
        continue
    if not insec: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,6}\*/', l) and cur: cnt[cur] += 1
src = {}
for f in ('pf_kernels.cuh', 'pf_math.cuh', 'pf_tail_dev.cuh', 'pf_tail.cuh'):
    src[f] = open(os.path.join(ROOT, 'pf_monocular_pose_estimator_b200', 'csrc', f)).read().split('\n')
def region(f, ln):
    if f not in src: return f
    for l in range(min(ln, len(src[f])) - 1, -1, -1):
        m = re.match(r'^(?:PF_HD|__device__|__global__|static|template).*?\b(pf_\w+)\s*\(', src[f][l])
        if m: return f + ':' + m.group(1)
    return f
agg = collections.Counter()
for (f, ln), c in cnt.items(): agg[region(f, ln)] += c
print("instructions:", sum(cnt.values()), "=", sum(cnt.values()) * 16, "bytes")
for k, c in agg.most_common(int(sys.argv[1]) if len(sys.argv) > 1 else 25): print('%6d  %s' % (c, k))

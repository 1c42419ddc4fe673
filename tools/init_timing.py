"""Wall time of the whole initialisation (device vote + device checkCorrespondences bodies + the reference's host
tail, tests/pose_estimator_init.cpp) for a few outlier counts; needs a GPU."""
import os, subprocess, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import capi, scenes as S
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pf_monocular_pose_estimator_b200")
capi.lib()
d = tempfile.mkdtemp()
exe = os.path.join(d, "init")
subprocess.check_call(["g++", "-std=c++11", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "pose_estimator_init.cpp"),
                       "-o", exe, "-L", PKG, "-lpfmpe", "-Wl,-rpath," + PKG])
for n_out in (0, 10, 20, 40):
    sc = S.Scene(n_outliers=n_out)
    p = os.path.join(d, "s.txt")
    with open(p, "w") as f:
        f.write("2000 5 3 1\n" + "\n".join("%r %r %r" % tuple(map(float, m)) for m in sc.markers) + "\n")
        f.write(" ".join(repr(float(v)) for v in sc.K.reshape(-1)) + "\n")
        b = sc.blobs(7)
        f.write("%r %d\n" % (sc.time(7), b.shape[0]) + "\n".join("%r %r" % (float(q[0]), float(q[1])) for q in b) + "\n0\n")
    r = subprocess.run([exe, "init", p, "time"], capture_output=True, text=True)
    print(r.stdout.splitlines()[0][:60], "|", " ".join(l.split(":")[1].strip() for l in r.stderr.splitlines()))

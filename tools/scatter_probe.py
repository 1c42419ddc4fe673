"""One GPU, W logical shards of one filter: device time of every shard's rebalance scatter per frame, with the number
of slots it produced and the number of runs (distinct sources) among them."""
import sys

import numpy as np
import torch

from pf_monocular_pose_estimator_b200 import capi, scenes, sharded
from tests.host_loop import HostTracker

W = int(sys.argv[1]) if len(sys.argv) > 1 else 8
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 25
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 10
log = []
orig = capi.Handle.shard_scatter


def timed(self, first, count):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    orig(self, first, count)
    b.record()
    log.append((int(count), a, b))


capi.Handle.shard_scatter = timed
sc = scenes.Scene(0)
sh = sharded.ShardedFilter(N, sharded.Loopback(W), n_leds=5, max_blobs=sc.blobs(0).shape[0] + 4, device=0)
sh.configure(sc.markers, sc.K, sc.params)
tb = HostTracker(sh, sc, init_particles=False)
sh.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
for f in range(2, 2 + frames):
    log.clear()
    rb = tb.frame(f)
    torch.cuda.synchronize()
    rows = ["%d:%.3f" % (c, a.elapsed_time(b)) for c, a, b in log]
    print("frame", f, "max_count", rb["max_count"], "slots:ms per shard", " ".join(rows), flush=True)

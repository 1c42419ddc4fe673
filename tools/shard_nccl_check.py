"""torchrun check of the sharded filter over NCCL: every rank runs (a) its shard of the filter through
sharded.TorchDist and (b) the whole unsharded filter on its own GPU, and compares them bit for bit
(record, its slice of the weights / propagated particles / rebalanced resampled set).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/shard_nccl_check.py [--particles 262144] [--frames 6]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from pf_monocular_pose_estimator_b200 import capi, scenes, sharded
from tests.host_loop import HostTracker

FIELDS = ["status", "iterations", "best_iteration", "resampled", "most_likely_particle_idx", "most_likely_pose_idx",
          "max_count", "highest_prob", "weight_sum"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=1 << 18)
    ap.add_argument("--frames", type=int, default=6)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N = a.particles // world * world
    n_loc = N // world
    sc = scenes.Scene(0)
    D = sc.blobs(0).shape[0] + 4
    ref = capi.Handle(N, n_trackers=1, n_leds=5, max_blobs=D, device=local)
    ref.configure(sc.markers, sc.K, sc.params)
    sh = sharded.ShardedFilter(N, sharded.TorchDist(), n_leds=5, max_blobs=D, device=local)
    sh.configure(sc.markers, sc.K, sc.params)
    ta = HostTracker(ref, sc)
    tb = HostTracker(sh, sc, init_particles=False)
    sh.set_particles([np.repeat(tb.cur[None], n_loc, axis=0)])
    bad, t_sh = 0, []
    for f in range(2, 2 + a.frames):
        ra = ta.frame(f)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        rb = tb.frame(f)
        torch.cuda.synchronize()
        t_sh.append(time.perf_counter() - t0)
        for k in FIELDS:
            bad += int(ra[k] != rb[k])
        bad += int(not (ra["pose"] == rb["pose"]).all()) + int(not (ra["correspondences"] == rb["correspondences"]).all())
        sl = slice(rank * n_loc, (rank + 1) * n_loc)
        bad += int(not (ref.weights(False)[sl] == sh.weights(False)).all())
        bad += int(not (ref.particles(0)[sl] == sh.particles(0)).all())
        if ra["resampled"]:
            bad += int(not (ref.particles(1)[sl] == sh.particles(1)).all())
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print(json.dumps(dict(check="shard_nccl", world=world, particles=N, frames=a.frames, mismatches=int(t.item()),
                              ms_per_frame_host_loop=round(1e3 * float(np.median(t_sh)), 3))))
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()

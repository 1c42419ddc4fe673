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
    ap.add_argument("--no-compare", action="store_true", help="timing only (large filters: the getters copy everything to the host)")
    ap.add_argument("--shards-per-gpu", dest="virtual", type=int, default=1, help="shards per GPU (block-cyclic placement), world * virtual <= 32")
    ap.add_argument("--stages", action="store_true", help="per-stage wall times (synchronises after every stage)")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    V = a.virtual
    N = a.particles // (world * V) * (world * V)
    n_loc = N // (world * V)
    sc = scenes.Scene(0)
    D = sc.blobs(0).shape[0] + 4
    cmp_ = not a.no_compare
    if cmp_:
        ref = capi.Handle(N, n_trackers=1, n_leds=5, max_blobs=D, device=local)
        ref.configure(sc.markers, sc.K, sc.params)
        ta = HostTracker(ref, sc, init_particles=False)
        ref.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
    sh = sharded.ShardedFilter(N, sharded.TorchDist(virtual=V) if world > 1 else sharded.Loopback(V), n_leds=5, max_blobs=D, device=local)
    sh.configure(sc.markers, sc.K, sc.params)
    tb = HostTracker(sh, sc, init_particles=False)
    sh.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
    bad, t_sh = 0, []
    if a.stages:
        sh.stage_s = {}
    for f in range(2, 2 + a.frames):
        if cmp_:
            ra = ta.frame(f)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        rb = tb.frame(f)
        torch.cuda.synchronize()
        t_sh.append(time.perf_counter() - t0)
        if a.stages and rank == 0:
            w = sh.weights(False)
            nz = np.flatnonzero(w > 0)
            print("frame", f, {k: round(1e3 * v, 3) for k, v in sh.stage_s.items()}, "max_count", rb["max_count"], "iters", rb["iterations"],
                  "nonzero", nz.size, "span", (int(nz[0]), int(nz[-1])) if nz.size else None, "wmax", float(w.max()), flush=True)
            sh.stage_s = {}
        if not cmp_:
            continue
        for k in FIELDS:
            bad += int(ra[k] != rb[k])
        bad += int(not (ra["pose"] == rb["pose"]).all()) + int(not (ra["correspondences"] == rb["correspondences"]).all())
        # this process hosts the shards rank, rank + world, ... (block-cyclic for V > 1): their slices of the unsharded filter
        ids = sh.comm.local
        sl = np.concatenate([np.arange(q * n_loc, (q + 1) * n_loc) for q in ids])
        bad += int(not (ref.weights(False)[sl] == sh.weights(False)).all())
        bad += int(not (ref.particles(0)[sl] == sh.particles(0)).all())
        if ra["resampled"]:
            bad += int(not (ref.particles(1)[sl] == sh.particles(1)).all())
    t = torch.tensor([bad], device="cuda")
    if world > 1:
        dist.all_reduce(t)
    if rank == 0:
        print(json.dumps(dict(check="shard_nccl", virtual=V, loop=sh.loop, rebalance=sh.rebalance, partials=sh.partial_exchange, p2p_error=sh.p2p_error, world=world, particles=N, frames=a.frames, mismatches=int(t.item()),
                              ms_per_frame_host_loop=round(1e3 * float(np.median(t_sh)), 3),
                              all_ms=[round(1e3 * v, 2) for v in t_sh])))
    if world > 1:
        dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()

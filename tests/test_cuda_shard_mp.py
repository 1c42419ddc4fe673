"""The sharded filter across REAL processes and GPUs: two ranks, one GPU each, CUDA-IPC peer
mailboxes + peer-memory rebalance, the C frame loop (pfmpe_shard_frame) -- bit-identical to the unsharded filter that
every rank also runs on its own GPU.  Skipped on a box with fewer than two GPUs (tools/shard_nccl_check.py is the
torchrun form of the same check for 2 / 4 / 8 ranks)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_global, frames, virtual, q, mean_cov=False):
    import torch
    import torch.distributed as dist
    from pf_monocular_pose_estimator_b200 import capi, scenes as S, sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        sc = S.Scene(0)
        D = sc.blobs(0).shape[0]
        flags = capi.FLAG_MEAN_COV if mean_cov else 0
        ref = capi.Handle(n_global, n_trackers=1, n_leds=5, max_blobs=D + 4, device=rank, flags=flags)
        ref.configure(sc.markers, sc.K, sc.params)
        ref.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
        sh = sharded.ShardedFilter(n_global, sharded.TorchDist(virtual=virtual), n_leds=5, max_blobs=D + 4, device=rank, rebalance="p2p", loop="c", flags=flags)
        sh.configure(sc.markers, sc.K, sc.params)
        sh.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
        ok, why = True, ""
        for f in range(2, 2 + frames):
            a = ref.track_frame_host(sc.time(f), sc.blobs(f), sc.seed(f))
            b = sh.track_frame_host(sc.time(f), sc.blobs(f), sc.seed(f))
            for k in ("status", "iterations", "most_likely_pose_idx", "max_count", "highest_prob", "weight_sum", "gn_iterations"):
                if a[k] != b[k]:
                    ok, why = False, "frame %d: %s %r != %r" % (f, k, a[k], b[k])
            if not ((a["pose"] == b["pose"]).all() and (a["cov"] == b["cov"]).all()):
                ok, why = False, "frame %d: pose / covariance differ" % f
            mine = torch.tensor([c - (1 << 64) if c >= (1 << 63) else c for c in sh.state_checksums()], dtype=torch.int64, device="cuda")
            dist.all_reduce(mine)                                     # int64 addition wraps: the sum mod 2^64
            tot = [int(v) & 0xFFFFFFFFFFFFFFFF for v in mine.cpu()]
            if tot != ref.state_checksum():
                ok, why = False, "frame %d: state checksums differ %r %r" % (f, tot, ref.state_checksum())
        if mean_cov:      # K7 across the ranks: one more PF update from the tracked state, mean pose + covariance of the whole filter
            f = 2 + frames
            I4 = np.eye(4)
            ua = ref.update(sc.blobs(f), a["pose"], a["pose"], I4, I4, 2, 0.02, sc.seed(f))
            ub = sh.update(sc.blobs(f), a["pose"], a["pose"], I4, I4, 2, 0.02, sc.seed(f))
            if ua["weight_sum"] != ub["weight_sum"] or ua["most_likely_pose_idx"] != ub["most_likely_pose_idx"]:
                ok, why = False, "update: PF results differ"
            if not (np.abs(ub["mean_pose"] - ua["mean_pose"]).max() < 1e-10 and np.allclose(ub["mean_cov"], ua["mean_cov"], rtol=1e-8, atol=1e-16)
                    and np.abs(ua["mean_cov"]).max() > 0):
                ok, why = False, "update: mean pose / covariance differ"
        q.put((rank, ok, why, sh.loop, sh.rebalance))
        ref.close()
        sh.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_global,virtual,mean_cov", [(200_000, 1, True), (1 << 22, 1, False), (1 << 20, 4, False), (1 << 18, 2, True)])
def test_two_processes_two_gpus_match_the_unsharded_filter(n_global, virtual, mean_cov):
    """virtual = 4: block-cyclic placement, 8 shards, process p hosts the shards p, p + 2, p + 4, p + 6.
    mean_cov: PFMPE_FLAG_MEAN_COV on every handle -- K7's two exchanges ride on every frame, and the weighted mean pose /
    covariance of the whole filter is compared at the end."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_global, 6, virtual, q, mean_cov)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("a rank hung")
        assert p.exitcode == 0
    res = sorted(q.get(timeout=5) for _ in range(2))
    for rank, ok, why, loop, reb in res:
        assert ok, (rank, why)
        assert loop == "c" and reb == "p2p"

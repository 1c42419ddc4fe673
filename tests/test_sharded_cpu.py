"""Host logic of the sharded filter (pf_monocular_pose_estimator_b200/sharded.py): ordered aggregates,
slot partition, pass reduction; and the torch.distributed communicator on 2 gloo ranks (CPU)."""
import os
import socket

import numpy as np
import pytest

from pf_monocular_pose_estimator_b200 import sharded


def test_unord_inverts_the_ordered_mapping():
    for v in [0.0, 1.0, -1.0, 4.999, 25.0, -3.5e-9, 1e30]:
        u = np.array([v], dtype=np.float32).view(np.uint32)[0]
        o = int(u) ^ (0xFFFFFFFF if (u >> 31) else 0x80000000)
        assert sharded.unord(o) == np.float32(v)
    # order preserving
    vals = np.array([-7.0, -1e-3, 0.0, 1e-3, 2.5, 25.0], dtype=np.float32)
    keys = [int(u) ^ (0xFFFFFFFF if (u >> 31) else 0x80000000) for u in vals.view(np.uint32)]
    assert keys == sorted(keys)


def test_lower_aggregates_match_a_flat_prefix():
    rng = np.random.default_rng(3)
    w = rng.integers(-1000, 5000, size=4096)
    W, n = 8, 512

    def agg(x):
        c = np.cumsum(x)
        return (int(c[-1]), int(c.max()), int(c.min()))

    aggs = [agg(w[r * n:(r + 1) * n]) for r in range(W)]
    lower, total = sharded.lower_aggregates(aggs)
    assert lower[0] == sharded.SM3_IDENTITY
    for r in range(1, W):
        assert lower[r] == agg(w[:r * n])
    assert total == agg(w)


def test_slot_partition_tiles_and_conserves():
    n_loc, W = 1000, 4
    first, end = [0, 2500, 2500, 3100], [2500, 2500, 3100, 4000]
    send = sharded.slot_partition(first, end, n_loc, W)
    sharded.check_tiling(first, end, n_loc * W)
    assert send[0] == [1000, 1000, 500, 0] and send[1] == [0, 0, 0, 0]
    assert send[2] == [0, 0, 500, 100] and send[3] == [0, 0, 0, 900]
    assert [sum(send[r][d] for r in range(W)) for d in range(W)] == [n_loc] * W
    # everything owned by one rank; trailing ranks start at n_global
    first, end = [0, 0, 4000, 4000], [0, 4000, 4000, 4000]
    sharded.check_tiling(first, end, 4000)
    send = sharded.slot_partition(first, end, n_loc, W)
    assert send[1] == [1000] * 4 and sum(map(sum, send)) == 4000
    with pytest.raises(RuntimeError):
        sharded.check_tiling([0, 2000], [1000, 4000], 4000)
    with pytest.raises(RuntimeError):
        sharded.check_tiling([0, 1000], [1000, 3000], 4000)


def test_reduce_pass_uses_unsigned_keys_and_exact_sums():
    p = np.zeros((3, 8), dtype=np.int64)
    keys = np.array([(0xC1C80000 << 32) | 0xFFFFFFF0, (0xC1C80000 << 32) | 0xFFFFFF00, (0x3F800000 << 32) | 5], dtype=np.uint64)
    p[:, 0] = keys.view(np.int64)
    p[:, 1] = keys[::-1].view(np.int64)
    p[:, 2] = [1 << 40, -(1 << 39), 7]
    p[:, 3:6] = [[1, 2, 3], [4, 5, 6], [7, 8, 9]]
    r = sharded.reduce_pass(p)
    assert r["maxkey"] == int(keys[0]) and r["minkey_inv"] == int(keys[0])     # first index wins the tie on the value
    assert r["sum"] == (1 << 40) - (1 << 39) + 7
    assert r["aggs"][1] == (4, 5, 6)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = sharded.TorchDist()
        part = torch.full((1, 8), rank + 1, dtype=torch.int64)
        g = comm.gather(part)
        ok = g.shape == (world, 8) and (g[:, 0] == np.arange(1, world + 1)).all()
        # rank 0 produced slots [0, 1500), rank 1 [1500, 2000): n_loc = 1000
        first, end, n_loc = [0, 1500], [1500, 2000], 1000
        send = sharded.slot_partition(first, end, n_loc, world)
        cnt = end[rank] - first[rank]
        rows = (torch.arange(first[rank], end[rank], dtype=torch.float32)[:, None] * torch.ones((1, 12))).contiguous()
        assert rows.shape == (cnt, 12)
        new = torch.empty((n_loc, 12))
        comm.exchange([rows], send, [new])
        ok = ok and bool((new[:, 0] == torch.arange(rank * n_loc, (rank + 1) * n_loc)).all())
        b = comm.bcast(np.arange(5.0) * (rank + 1), 1)
        ok = ok and (b == np.arange(5.0) * 2).all()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_torchdist_communicator_on_two_gloo_ranks():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got = dict(q.get(timeout=10) for _ in range(2))
    assert got == {0: True, 1: True}

"""One particle filter sharded over several GPUs (BASELINE configs[4]).

The reference runs ONE filter per tracked object (pose_estimator.cpp:471-735).  For the giant
single-filter case its particle set is split into `world` contiguous, equal shards, one per GPU /
process.  The loop is cut at its global reductions; every stage is one launch of the sm_100a frame
kernel through the C ABI (pfmpe_shard_*, include/pfmpe.h) and this module owns the exchange steps:

  score pass   -> all_gather of one 64-byte partial per rank  (max / min / weight sum / prefix aggregate:
                  the "NCCL weight-sum allreduce" of the north star; a gather keeps the per-rank
                  aggregates that the ordered prefix needs, 64 B x world)
  decision     -> do/while exit, kept iteration, accept test: host scalars, same on every rank (:608-633)
  resampling   -> every rank resamples its own sources against GLOBAL strata (global slot numbers, the
                  global RNG stream); all_gather of (count key, first owned slot, end slot)
  rebalance    -> every rank again owns n_global / world particles in global slot order ("global
                  resample"): slots that stay on their rank are copied on the device, the others
                  travel in one all_to_all
  winner       -> first most-resampled particle across ranks (:685-686); its owner computes pose and
                  correspondences, broadcast to the others

Because RNG addressing, strata, keys and tie rules all use global particle / slot numbers, the sharded
filter reproduces the single-GPU filter bit for bit (tests/test_cuda_shard.py checks exactly that with
logical shards on one GPU and with NCCL ranks).

Two communicators: `Loopback` (all shards in this process, exchange = device-local copies; used by
the single-GPU tests and as the W=1 case) and `TorchDist` (one shard per process, torch.distributed).
The host logic (`reduce_pass`, `slot_partition`, `lower_aggregates`) is pure numpy/python and is
covered by the CPU tests, including a world_size-2 gloo run of `TorchDist`.
"""
import struct

import numpy as np

from . import capi

Q30 = 1 << 30
SM_NEG = -(1 << 62)      # PF_SM_NEG / PF_SM_POS of pf_math.cuh ("no prefix yet")
SM_POS = 1 << 62
MAX_ITER = 80            # PF_MAX_ITER (:616)


# ---------------------------------------------------------------------------------------
# host logic (no device, no torch)
# ---------------------------------------------------------------------------------------
def unord(o):
    """Inverse of the ordered-uint mapping of an f32 (pf_ord / pf_unord in pf_kernels.cuh)."""
    o &= 0xFFFFFFFF
    u = (o ^ 0x80000000) if (o & 0x80000000) else (~o & 0xFFFFFFFF)
    return struct.unpack("<f", struct.pack("<I", u))[0]


def sm3_combine(a, b):
    """Ordered combination of (sum, max prefix, min prefix) aggregates (pf_sm3_combine)."""
    tx = SM_NEG if b[1] <= SM_NEG else a[0] + b[1]
    tn = SM_POS if b[2] >= SM_POS else a[0] + b[2]
    return (a[0] + b[0], max(a[1], tx), min(a[2], tn))


SM3_IDENTITY = (0, SM_NEG, SM_POS)


def reduce_pass(partials):
    """partials: (world, 8) int64 of one score pass -> dict(maxkey, minkey_inv, sum, aggs)."""
    u = np.ascontiguousarray(partials[:, :2]).view(np.uint64)
    return dict(maxkey=int(u[:, 0].max()), minkey_inv=int(u[:, 1].max()),
                sum=int(sum(int(v) for v in partials[:, 2])),
                aggs=[tuple(int(v) for v in row[3:6]) for row in partials])


def lower_aggregates(aggs):
    """Exclusive ordered prefix over the ranks: (per-rank aggregate of all lower ranks, total)."""
    lower, run = [], SM3_IDENTITY
    for a in aggs:
        lower.append(run)
        run = sm3_combine(run, a)
    return lower, run


def slot_partition(first, end, n_loc, world):
    """send[r][d] = number of output slots of [first[r], end[r]) that fall into shard d's range
    [d * n_loc, (d + 1) * n_loc).  The slot ranges of the ranks tile [0, world * n_loc) in rank order."""
    send = [[0] * world for _ in range(world)]
    for r in range(world):
        for d in range(world):
            lo, hi = max(int(first[r]), d * n_loc), min(int(end[r]), (d + 1) * n_loc)
            send[r][d] = max(0, hi - lo)
    return send


def check_tiling(first, end, n_global):
    pos = 0
    for a, b in zip(first, end):
        if int(b) < int(a) or (int(b) > int(a) and int(a) != pos):
            raise RuntimeError("resampled slot ranges do not tile the filter: %r %r" % (list(first), list(end)))
        if int(b) > int(a):
            pos = int(b)
    if pos != n_global:
        raise RuntimeError("resampled slot ranges cover %d of %d slots" % (pos, n_global))


# ---------------------------------------------------------------------------------------
# communicators
# ---------------------------------------------------------------------------------------
class Loopback:
    """All `world` shards live in this process on one device; exchange steps are local copies."""

    def __init__(self, world):
        self.world = world
        self.local = list(range(world))

    def gather(self, partials):
        return partials.cpu().numpy()

    def exchange(self, send_rows, send, outs):
        """send_rows[r]: rows leaving shard r ordered by destination; outs[d]: rows arriving, ordered by source."""
        import torch
        for d in self.local:
            parts = []
            for r in self.local:
                off, c = sum(send[r][:d]), send[r][d]
                if c:
                    parts.append(send_rows[r][off:off + c])
            if parts:
                torch.cat(parts, out=outs[d])

    def bcast(self, arr, owner):
        return arr

    def barrier(self, dev):
        pass

    def all_ok(self, ok, dev):
        return bool(ok)

    def device_barrier(self, shards):
        pass                                                 # one stream: already ordered

    def peer_pointers(self, shards, dev):
        """Bases of every shard's next-frame particle planes (same process: plain device pointers)."""
        return [h.shard_local_ptr() for h in shards]

    def mail_pointers(self, shards, dev):
        return [h.shard_mail_create(self.world, r)[0] for r, h in enumerate(shards)]


class TorchDist:
    """One process per GPU; NCCL on GPUs, gloo in the CPU tests.

    virtual = V > 1: block-cyclic placement.  The filter is cut into procs * V shards and process p hosts the shards
    {p, p + procs, p + 2 procs, ...} (V handles on its GPU, one stream), so that every GPU holds V blocks spread over
    the global particle order: when the weight concentrates in one stretch of the order -- the usual state of a
    converged filter -- the resampling work and the cross-GPU traffic of the rebalance spread over the GPUs instead of
    falling on one.  Needs the C frame loop (peer mailboxes + peer-memory rebalance); procs * V <= 32."""

    def __init__(self, group=None, virtual=1):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.procs = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.virtual = int(virtual)
        self.world = self.procs * self.virtual
        self.local = [self.rank + j * self.procs for j in range(self.virtual)]

    def _need_single(self, what):
        if self.virtual != 1:
            raise RuntimeError("%s: block-cyclic placement (virtual > 1) runs through the C frame loop only" % what)

    def gather(self, partials):
        import torch
        self._need_single("all_gather of the partials")
        out = torch.empty((self.world, partials.shape[1]), dtype=partials.dtype, device=partials.device)
        self.dist.all_gather_into_tensor(out, partials.reshape(-1, partials.shape[1]), group=self.group)
        return out.cpu().numpy()

    def exchange(self, send_rows, send, outs):
        self._need_single("row exchange")
        r = self.rank
        recv = [send[q][r] for q in range(self.world)]
        self.dist.all_to_all_single(outs[0], send_rows[0], output_split_sizes=recv, input_split_sizes=list(send[r]),
                                    group=self.group)

    def barrier(self, dev):
        """Stream-ordered: every rank's earlier work on its current stream precedes everybody's later work."""
        import torch
        if not hasattr(self, "_one"):
            self._one = torch.zeros(1, dtype=torch.int32, device=dev)
        self.dist.all_reduce(self._one, group=self.group)

    def device_barrier(self, shards):
        shards[0].shard_barrier()                            # signal + wait kernels through the peer mailboxes

    def all_ok(self, ok, dev):
        import torch
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN, group=self.group)
        return bool(int(t.item()))

    def _all_handles(self, mine, dev):
        """mine: V 64-byte CUDA IPC handles of this process's shards -> (procs, V, 64) uint8 of everybody's."""
        import torch
        t = torch.tensor([list(hd) for hd in mine], dtype=torch.uint8, device=dev).reshape(1, self.virtual, 64)
        allh = torch.empty((self.procs, self.virtual, 64), dtype=torch.uint8, device=dev)
        self.dist.all_gather_into_tensor(allh, t, group=self.group)
        return allh.cpu().numpy()

    def peer_pointers(self, shards, dev):
        """CUDA IPC: every process publishes the buffers of its shards and opens the others' (NVLink peer access).
        Returns the pointers in shard order (shard q lives on process q % procs as its local shard q // procs)."""
        allh = self._all_handles([h.shard_ipc_export() for h in shards], dev)
        return [shards[q // self.procs].shard_local_ptr() if q % self.procs == self.rank
                else shards[0].shard_ipc_open(allh[q % self.procs, q // self.procs].tobytes()) for q in range(self.world)]

    def mail_pointers(self, shards, dev):
        made = [h.shard_mail_create(self.world, q) for h, q in zip(shards, self.local)]
        allh = self._all_handles([hd for _, hd in made], dev)
        return [made[q // self.procs][0] if q % self.procs == self.rank
                else shards[0].shard_ipc_open(allh[q % self.procs, q // self.procs].tobytes()) for q in range(self.world)]

    def bcast(self, arr, owner):
        import torch
        self._need_single("broadcast of the winner")
        dev = "cuda" if self.dist.get_backend(self.group) == "nccl" else "cpu"
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64)).to(dev)
        src = owner if self.group is None else self.dist.get_global_rank(self.group, owner)
        self.dist.broadcast(t, src=src, group=self.group)
        return t.cpu().numpy()


# ---------------------------------------------------------------------------------------
# the sharded filter
# ---------------------------------------------------------------------------------------
class ShardedFilter:
    """Drop-in for capi.Handle.update() on one tracked object whose particles are sharded.

    n_global must be a multiple of comm.world.  With `Loopback` all shards are handles on `device`."""

    def __init__(self, n_global, comm, n_leds=5, max_blobs=32, device=0, rebalance="auto", mail=True, loop="auto", **kw):
        import torch
        if n_global % comm.world:
            raise ValueError("n_global must be a multiple of the number of shards")
        self.torch = torch
        self.comm, self.W = comm, comm.world
        self.N, self.n_loc, self.L, self.T = n_global, n_global // comm.world, n_leds, 1
        self.dev = torch.device("cuda", device)
        torch.cuda.set_device(self.dev)
        self.shards = []
        for r in comm.local:
            h = capi.Handle(self.n_loc, n_trackers=1, n_leds=n_leds, max_blobs=max_blobs, device=device, **kw)
            h.shard_configure(r * self.n_loc, n_global)
            h.set_stream(torch.cuda.current_stream(self.dev).cuda_stream)
            self.shards.append(h)
        self.partials = torch.zeros((len(self.shards), 32), dtype=torch.int64, device=self.dev)
        self._bufs = {}
        # rebalance of the resampled set: "p2p" = one scatter kernel with NVLink peer stores (CUDA IPC),
        # "exchange" = row export + all_to_all + import; "auto" tries p2p and falls back for every rank
        self.rebalance, self.p2p_error = "exchange", None
        if rebalance in ("auto", "p2p"):
            ok = True
            try:
                ptrs = comm.peer_pointers(self.shards, self.dev)
                for h in self.shards:
                    h.shard_set_peers(ptrs)
            except capi.PfmpeError as e:
                ok, self.p2p_error = False, str(e)
            if comm.all_ok(ok, self.dev):
                self.rebalance = "p2p"
            elif rebalance == "p2p":
                raise RuntimeError("peer-memory rebalance unavailable: %s" % (self.p2p_error or "a peer rank failed"))
            else:
                for h in self.shards:                        # the peer tables also switch the kernels to lead[] gathers
                    h.shard_set_peers([])
        # exchange of the 64-byte partials: "mail" = peer stores from inside the stage kernels + a wait
        # kernel that hands the table to the host through pinned memory (needs peer access, like p2p);
        # otherwise all_gather + device->host copy
        self.partial_exchange = "all_gather"
        if self.rebalance == "p2p" and mail and self.W <= 32:
            ok = True
            try:
                ptrs = comm.mail_pointers(self.shards, self.dev)
                for h in self.shards:
                    h.shard_mail_set_peers(ptrs)
            except capi.PfmpeError as e:
                ok, self.p2p_error = False, str(e)
            if comm.all_ok(ok, self.dev):
                self.partial_exchange = "mail"
                comm.barrier(self.dev)                       # every mailbox is installed before the first message
        # host loop of the frame: "c" = pfmpe_shard_frame (the product path: decisions in C from the mailbox partials; needs
        # the peer mailboxes and the peer-memory rebalance), "python" = the stage-by-stage loop below (torch.distributed
        # fallbacks and the per-stage profiling aid)
        c_ok = self.partial_exchange == "mail" and self.rebalance == "p2p"
        if loop == "c" and not c_ok:
            raise RuntimeError("loop='c' needs the peer mailboxes and the peer-memory rebalance")
        self.loop = "c" if (c_ok and loop in ("auto", "c")) else "python"
        self._passes = 0         # score passes launched by the python loop (per shard)
        self.last = None
        self.stage_s = None      # set to {} to collect per-stage wall times (synchronises after every stage)

    def _buffer(self, kind, i, rows):
        """Grow-only row buffers (12 floats per row): no allocator traffic in the frame loop."""
        key = (kind, i)
        buf = self._bufs.get(key)
        if buf is None or buf.shape[0] < rows:
            buf = self.torch.empty((max(rows + rows // 4, 1024), 12), dtype=self.torch.float32, device=self.dev)
            self._bufs[key] = buf
        return buf[:rows]

    def _mark(self, name, t0):
        """Profiling aid: with stage_s set, wait for the device and book the time since t0 under `name`."""
        if self.stage_s is None:
            return 0.0
        import time
        self.torch.cuda.synchronize(self.dev)
        t1 = time.perf_counter()
        self.stage_s[name] = self.stage_s.get(name, 0.0) + (t1 - t0)
        return t1

    @property
    def passes(self):
        """Score passes launched so far (per shard)."""
        return self._passes + (self.shards[0].shard_passes() if self.shards else 0)

    def close(self):
        for h in self.shards:
            h.close()
        self.shards = []

    def state_checksums(self):
        """Position-keyed device checksums summed over the LOCAL shards (add the ranks' values mod 2^64 for the filter's)."""
        tot = [0, 0, 0]
        for h in self.shards:
            c = h.state_checksum()
            tot = [(a + b) & 0xFFFFFFFFFFFFFFFF for a, b in zip(tot, c)]
        return tot

    def track_frame_host(self, time_to_predict, image_points, seed):
        """One tracked frame from host buffers (predictPose -> sharded PF update -> Gauss-Newton), in C when the C loop is on."""
        if self.loop == "c":
            return capi.shard_track_frame_host(self.shards, time_to_predict, image_points, seed)
        raise RuntimeError("track_frame_host needs loop='c' (use tests.host_loop.HostTracker around update() otherwise)")

    def configure(self, markers, K, params):
        for h in self.shards:
            h.configure(markers, K, params)

    def set_particles(self, poses, tracker=0):
        """poses: (n_global, 4, 4), or (n_loc, 4, 4) per local shard if given as a list."""
        for i, r in enumerate(self.comm.local):
            p = poses[i] if isinstance(poses, list) else poses[r * self.n_loc:(r + 1) * self.n_loc]
            self.shards[i].set_particles(p)

    def track_init(self, pose0, t0, pose1, t1):
        """Every particle = pose1, filled on the device (no host-side particle array)."""
        for h in self.shards:
            h.track_init(pose0, t0, pose1, t1)

    # -- stages -------------------------------------------------------------------------
    def _gather(self):
        """Partials of the stage just launched on every rank -> (world, 8) int64 on the host."""
        if self.partial_exchange == "mail":
            return self.shards[0].shard_collect(self.W)[:, :8]
        return self.comm.gather(self.partials)[:, :8]

    def _score(self, it):
        t = self._mark("other", self._t)
        for i, h in enumerate(self.shards):
            h.shard_score(it, self.partials[i].data_ptr())
        self._passes += 1
        t = self._mark("score", t)
        r = reduce_pass(self._gather())
        self._t = self._mark("gather", t)
        return r

    def update(self, img, T_cur, T_pred, P, Cm, isi, dt, seed, tracker=0):
        torch = self.torch
        keep = []
        fi = capi.make_frame_in(img, T_cur, T_pred, P, Cm, isi, dt, seed, keep)
        if self.loop == "c" and self.stage_s is None:
            out = capi.shard_frame(self.shards, fi)
            out["sum_fx"] = int(round(out["weight_sum"] * Q30))
            self.last = out
            return out
        D = int(fi.n_image_points)
        import time
        self._t = time.perf_counter() if self.stage_s is not None else 0.0
        for h in self.shards:
            h.shard_begin_frame(fi)

        # ---- do { propagate, project, score } while (iter < 80 && max < L*min(5,D))  (:535-616)
        thr_exit = np.float32(self.L * min(D, 5))
        highest, best_it, it = np.float32(0.0), -1, 0
        kept = last = None
        while True:
            r = self._score(it)
            m = np.float32(unord(r["maxkey"] >> 32))
            last = r
            it += 1
            if m > highest:                                  # :608-614 strict >, from 0
                highest, best_it, kept = m, it - 1, r
            if it < MAX_ITER and m < thr_exit:
                continue
            break
        iters = it
        if best_it < 0:
            kept = last                                      # :619 nothing to restore
        elif best_it != iters - 1:
            kept = self._score(best_it)                      # regenerate the kept iteration (exact: RNG by address)

        S = kept["sum"]
        accept = S != 0 and highest > np.float32(self.L * min(D, 3))          # :633
        resampled = bool(accept and isi > 0)
        arg_max = (~kept["maxkey"]) & 0xFFFFFFFF
        arg_min = (~kept["minkey_inv"]) & 0xFFFFFFFF
        max_count = 0
        if resampled:
            lower, total = lower_aggregates(kept["aggs"])
            t = self._mark("other", self._t)
            for i, r_ in enumerate(self.comm.local):
                self.shards[i].shard_resample(S, iters, lower[r_], total, self.partials[i].data_ptr())
            t = self._mark("resample", t)
            part = self._gather()
            t = self._mark("gather", t)
            cntkey = int(np.ascontiguousarray(part[:, 0]).view(np.uint64).max())
            q = (~cntkey) & 0xFFFFFFFF                        # :686 first max of counterMeas
            max_count = cntkey >> 32
            first = part[:, 1]
            # a rank's slots end where the next rank's begin (its sources own nothing beyond); the last one ends at N
            end = np.concatenate([first[1:], [self.N]])
            if self.partial_exchange != "mail":
                assert (end == part[:, 2]).all(), (first, part[:, 2])
            check_tiling(first, end, self.N)
            send = slot_partition(first, end, self.n_loc, self.W)
            if self.rebalance == "p2p":
                for i, r_ in enumerate(self.comm.local):
                    self.shards[i].shard_scatter(int(first[r_]), int(end[r_] - first[r_]))
                # all scatters land before anyone's next score pass: a barrier on the ranks' streams
                if self.partial_exchange == "mail":
                    self.comm.device_barrier(self.shards)
                else:
                    self.comm.barrier(self.dev)
                self._t = self._mark("scatter", t)
            else:
                self._rebalance_exchange(send, first, end, t)
        else:
            q = arg_min if S < 0 else arg_max                 # :714 on the normalised weights
        return self._finish(q, S, accept, resampled, iters, best_it, highest, arg_max, max_count)

    def _rebalance_exchange(self, send, first, end, t):
        # slots that stay on their rank go plane to plane on the device; only the rest is exchanged
        send_x = [[0 if d == r_ else c for d, c in enumerate(row)] for r_, row in enumerate(send)]
        rows = []
        for i, r_ in enumerate(self.comm.local):
            lo, hi, f0, e0 = r_ * self.n_loc, (r_ + 1) * self.n_loc, int(first[r_]), int(end[r_])
            below = (f0, max(f0, min(e0, lo)))                 # to lower ranks
            above = (min(e0, max(f0, hi)), e0)                 # to higher ranks
            nb, na = below[1] - below[0], above[1] - above[0]
            assert nb + na == sum(send_x[r_]) and nb + na + send[r_][r_] == e0 - f0
            buf = self._buffer("send", i, nb + na)
            self.shards[i].shard_export(f0, below[0], nb, buf.data_ptr())
            self.shards[i].shard_export(f0, above[0], na, buf.data_ptr() + 48 * nb)
            self.shards[i].shard_keep(f0, below[1], send[r_][r_], below[1] - lo if send[r_][r_] else 0)
            rows.append(buf)
        t = self._mark("export", t)
        outs = [self._buffer("recv", i, sum(send_x[q][r_] for q in range(self.W))) for i, r_ in enumerate(self.comm.local)]
        self.comm.exchange(rows, send_x, outs)
        t = self._mark("exchange", t)
        for i, r_ in enumerate(self.comm.local):
            a = sum(send_x[q][r_] for q in range(r_))
            b = sum(send_x[q][r_] for q in range(r_ + 1, self.W))
            if a + send[r_][r_] + b != self.n_loc:
                raise RuntimeError("rebalance delivers %d rows to shard %d, expected %d" % (a + send[r_][r_] + b, r_, self.n_loc))
            self.shards[i].shard_import(outs[i].data_ptr(), 0, a)
            self.shards[i].shard_import(outs[i].data_ptr() + 48 * a, self.n_loc - b, b)
        self._t = self._mark("import", t)

    def _finish(self, q, S, accept, resampled, iters, best_it, highest, arg_max, max_count):
        # ---- winner pose + correspondences from its owner
        owner = q // self.n_loc
        packed = np.zeros(16 + 1 + 2 * capi.MAX_LED)
        if self.partial_exchange == "mail":
            # the owner's kernel sends pose + correspondences to every rank's mailbox (peer stores)
            for i, r_ in enumerate(self.comm.local):
                if r_ == owner:
                    self.shards[i].shard_winner_publish(q - owner * self.n_loc)
                else:
                    self.shards[i].shard_advance()
            row = np.ascontiguousarray(self.shards[0].shard_collect(self.W, only_rank=owner)[owner])
            pose = np.eye(4)
            pose[:3, :] = row[:6].view(np.float32).astype(np.float64).reshape(3, 4)
            nc = int(row[6 + capi.MAX_LED])
            packed[:16] = capi.colmajor(pose)
            packed[16] = nc
            packed[17:17 + 2 * nc] = row[6:6 + capi.MAX_LED].view(np.uint32)[:2 * nc]
        else:
            if owner in self.comm.local:
                w = self.shards[self.comm.local.index(owner)].shard_winner(q - owner * self.n_loc)
                packed[:16] = capi.colmajor(w["winner_pose"])
                c = w["correspondences"].reshape(-1)
                packed[16] = c.size // 2
                packed[17:17 + c.size] = c
            packed = self.comm.bcast(packed, owner)
        self._t = self._mark("winner", self._t)
        nc = int(packed[16])
        out = dict(status=1 if accept else 4, iterations=iters, best_iteration=best_it, resampled=int(resampled),
                   highest_prob=float(highest), weight_sum=S / Q30,
                   most_likely_particle_idx=(arg_max if best_it >= 0 else -1) if accept else q,
                   most_likely_pose_idx=q if resampled else -1, max_count=max_count,
                   correspondences=packed[17:17 + 2 * nc].astype(np.uint32).reshape(-1, 2),
                   winner_pose=capi.from_colmajor(packed[:16]), sum_fx=S)
        self.last = out
        return out

    # -- getters (local shards; concatenate across ranks yourself in distributed runs) ----
    def particles(self, which=0, tracker=0):
        return np.concatenate([h.particles(which) for h in self.shards])

    def weights(self, normalised=False, tracker=0):
        w = np.concatenate([h.weights(False) for h in self.shards])
        if normalised and self.last and self.last["sum_fx"] != 0:
            w = w / (self.last["sum_fx"] / Q30)
        return w

    def sync(self):
        for h in self.shards:
            h.sync()

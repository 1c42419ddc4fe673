#!/usr/bin/env python
"""bench.py -- particle-marker evaluations/sec (PME/s) of the PF tracking core on B200.

A "step" is one PF frame update (propagate -> project -> score loop, normalise, stratified
resample, winner, FP64 tail) of the workload's filter(s) on synthetic detections.

  N=1  default workload = BASELINE.json configs[1]: README 5-LED scene, 1e5 particles, one tracker,
       1000-frame synthetic trajectory ("cfg2").  Other workloads: --workload cfg1|cfg3|cfg4|cfg5s.
  N>1  default workload = configs[3]: 4096 independent trackers x 4096 particles sharded over the
       ranks (no data-path collective; torch.distributed/NCCL only for the barrier and the max-time).

value : PME/s with every input resident in HBM when the timed region starts (device-resident
        tracking, one launch per step, device time from CUDA events on the launching stream).
e2e   : the same metric through the reference-facing per-frame C-ABI call with HOST buffers: per
        step the detections / poses go host->device, the result record comes back device->host and
        the reference's host tail (Gauss-Newton + predictPose, FP64) runs between two updates.
--impl reference : the reference's CPU algorithm (oracle port, see oracle/pf_oracle.h) on the
        host cores, same workload / metric / unit.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from pf_monocular_pose_estimator_b200 import scenes as S  # noqa: E402

L_LED = 5
WORKLOADS = {
    # name: (particles, trackers, outliers, default frames)
    "cfg1": (100, 1, 10, 1000),
    "cfg2": (100_000, 1, 10, 1000),
    "cfg3": (1_000_000, 1, 200, 200),
    "cfg4": (4096, 4096, 10, 200),
    "cfg5s": (1 << 26, 1, 10, 50),
}
WORKLOAD_DESC = {
    "cfg1": "BASELINE configs[0]: 5-LED README scene, 100 particles, 1 tracker, 15 blobs",
    "cfg2": "BASELINE configs[1]: 5-LED README scene, 1e5 particles, 1 tracker, 15 blobs, synthetic trajectory",
    "cfg3": "BASELINE configs[2]: 5 LEDs + 200 spurious blobs, 1e6 particles, 1 tracker",
    "cfg4": "BASELINE configs[3]: 4096 independent trackers x 4096 particles, 15 blobs, sharded by tracker",
    "cfg5s": "BASELINE configs[4] on one GPU: single filter with 2^26 particles",
}


def alg_bytes_per_frame(n_particles, iters):
    """B_alg = N * (104 * I + 32) bytes per frame."""
    return n_particles * (104 * iters + 32)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self._stop = gpu, [], threading.Event()
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[1]) for r in self.rows)
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for i, nm in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][2]), "reasons": sorted(reasons),
                "samples": len(self.rows)}


def make_sequence(trackers, first_frame, n_frames, n_outliers):
    frames = list(range(first_frame, first_frame + n_frames))
    blobs, counts = S.blob_tensor_fast(trackers, frames, n_outliers=n_outliers)
    times = np.array([f * S.DT for f in frames])
    seeds = np.array([[S.pf_seed(f, int(t)) for f in frames] for t in trackers], dtype=np.uint32)
    return frames, blobs, counts, times, seeds


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference PF
# ---------------------------------------------------------------------------------------------
def cpu_run(workload, budget_s, threads):
    """Time the reference's CPU algorithm (oracle port) on a bounded sample of the workload."""
    from oracle import orc
    n_p, n_t, n_out, _ = WORKLOADS[workload]
    trackers = np.arange(min(n_t, threads))
    max_frames = 400
    frames, blobs, counts, times, seeds = make_sequence(trackers, 2, max_frames, n_out)
    sc0 = S.Scene()
    results = [None] * len(trackers)

    def work(i):
        t = int(trackers[i])
        ph = S.tracker_phase(t)
        trk = orc.OracleTracker(n_p, S.MARKERS_5, S.K_README, sc0.params, S.true_pose(0.0, ph), 0.0,
                                S.true_pose(S.DT, ph), S.DT, resampler=1)
        # one untimed frame (page-in), then time frames until the budget is used
        trk.frame(times[0], blobs[i, 0, : counts[i, 0]], int(seeds[i, 0]))
        pme, nfr, lat = 0, 0, []
        t0 = time.perf_counter()
        for f in range(1, max_frames):
            a = time.perf_counter()
            trk.frame(times[f], blobs[i, f, : counts[i, f]], int(seeds[i, f]))
            b = time.perf_counter()
            lat.append(b - a)
            pme += n_p * L_LED * trk.pf.iterations
            nfr += 1
            if b - t0 > budget_s:
                break
        results[i] = (pme, time.perf_counter() - t0, nfr, lat)
        trk.close()

    ths = [threading.Thread(target=work, args=(i,)) for i in range(len(trackers))]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    wall = time.perf_counter() - t0
    pme = sum(r[0] for r in results)
    lat = sorted(x for r in results for x in r[3])
    frames_done = sum(r[2] for r in results)
    return dict(value=pme / max(max(r[1] for r in results), 1e-9), cores=len(trackers), frames=frames_done,
                p50_ms=1e3 * lat[len(lat) // 2], wall_s=wall,
                sample="%d tracker(s) x %d frames of %s (%d particles each), O(N) resampler, double precision" %
                       (len(trackers), results[0][2], workload, n_p))


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
def gpu_run(args, rank, world, local_rank):
    import torch
    from pf_monocular_pose_estimator_b200 import capi
    from tests.host_loop import HostTracker

    workload = args.workload
    n_p, n_t_total, n_out, _ = WORKLOADS[workload]
    if n_t_total > 1:
        assert n_t_total % world == 0
        n_t = n_t_total // world
        trackers = np.arange(rank * n_t, (rank + 1) * n_t)
    else:
        n_t, trackers = 1, np.array([0])          # replicas only
    K, W = args.steps, args.warmup
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    d_max = L_LED + n_out
    frames, blobs, counts, times, seeds = make_sequence(trackers, 2, W + K, n_out)

    h = capi.Handle(n_p, n_trackers=n_t, n_leds=L_LED, max_blobs=d_max, device=local_rank)
    h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    for i, t in enumerate(trackers):
        ph = S.tracker_phase(int(t))
        h.track_init(S.true_pose(0.0, ph), 0.0, S.true_pose(S.DT, ph), S.DT, tracker=i)
    # inputs resident in HBM before the timed region
    d_blobs = torch.from_numpy(blobs).to(dev)
    d_counts = torch.from_numpy(counts).to(dev)
    d_times = torch.from_numpy(times).to(dev)
    d_seeds = torch.from_numpy(seeds.view(np.int32)).to(dev)
    h.track_attach(W + K, d_blobs.data_ptr(), d_counts.data_ptr(), d_times.data_ptr(), d_seeds.data_ptr())

    state_bytes = n_p * n_t * 108
    flush = None
    l2_note = "working set %.1f MB > L2 (126 MB): no flush needed" % (state_bytes / 1e6)
    if state_bytes < 400e6 and not args.no_flush:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        l2_note = "L2 flushed (256 MiB write, untimed) between steps; filter state is %.1f MB" % (state_bytes / 1e6)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize(dev)

    for f in range(W):
        if flush is not None:
            flush.zero_()
        h.track_run(f, 1)
    barrier()
    launches0 = h.launch_count()
    step_ms = []
    with ClockSampler(local_rank) as clk:
        t_wall0 = time.perf_counter()
        for f in range(W, W + K):
            if flush is not None:
                flush.zero_()
            h.track_run(f, 1)
            step_ms.append(h.last_launch_ms())      # CUDA events recorded on the launching stream
        barrier()
        t_wall = time.perf_counter() - t_wall0
    launches = h.launch_count() - launches0
    dev_ms = float(np.sum(step_ms))
    res = [h.track_results(W, K, tracker=i) for i in range(min(n_t, 8))]
    iters = np.array([[r["iterations"] for r in rr] for rr in res])
    ok = all(r["status"] == 1 for rr in res for r in rr)
    mean_iter = float(iters.mean())
    pme_rank = float(n_p) * L_LED * mean_iter * K * n_t
    alg_b = alg_bytes_per_frame(n_p, mean_iter) * n_t * K

    # same sequence, all K frames in ONE launch (state stays in L2 when it fits) -- explanatory only
    for i, t in enumerate(trackers):
        ph = S.tracker_phase(int(t))
        h.track_init(S.true_pose(0.0, ph), 0.0, S.true_pose(S.DT, ph), S.DT, tracker=i)
    h.track_run(0, W)
    h.sync()
    h.track_run(W, K)
    single_ms = h.last_launch_ms()

    if world > 1:
        tt = torch.tensor([dev_ms, single_ms, pme_rank, alg_b, float(launches), float(ok)], dtype=torch.float64, device=dev)
        mx = tt.clone(); torch.distributed.all_reduce(mx, op=torch.distributed.ReduceOp.MAX)
        sm = tt.clone(); torch.distributed.all_reduce(sm, op=torch.distributed.ReduceOp.SUM)
        mn = tt.clone(); torch.distributed.all_reduce(mn, op=torch.distributed.ReduceOp.MIN)
        dev_ms, single_ms = float(mx[0]), float(mx[1])
        pme_total, alg_total, launches, ok = float(sm[2]), float(sm[3]), int(sm[4]), bool(mn[5] > 0)
    else:
        pme_total, alg_total = pme_rank, alg_b

    # ---- e2e: per-frame C-ABI calls with host buffers + host FP64 tail (rank-local trackers: first one only
    # for multi-tracker workloads would under-count, so e2e uses update_batch over all local trackers)
    e2e = None
    if not args.no_e2e:
        e2e = e2e_run(h, capi, HostTracker, trackers, n_p, n_t, blobs, counts, times, seeds, W, min(K, args.e2e_steps), world, dev)

    peak, peak_src = peaks()
    achieved = alg_total / (dev_ms * 1e-3) / 1e9 / world      # per-GPU GB/s (max-time rank)
    line = {
        "metric": "particle_marker_evals_per_sec", "value": pme_total / (dev_ms * 1e-3), "unit": "PME/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True,
        "scaling": "strong" if n_t_total > 1 else "replicas", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload, "desc": WORKLOAD_DESC[workload], "particles": n_p, "trackers": n_t_total,
                   "leds": L_LED, "blobs": d_max, "mean_iterations": mean_iter, "l2": l2_note,
                   "launch": dict(zip(["grid_x", "grid_y", "block", "cooperative"], h.launch_info()))},
        "latency_p50_ms": float(np.median(step_ms)),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_source": peak_src,
                     "alg_bytes_per_particle_frame": alg_bytes_per_frame(1, mean_iter), "kernel": "pf_frames_kernel<5>"},
        "single_launch": {"value": pme_total / (single_ms * 1e-3), "ms_per_step": single_ms / K,
                          "note": "all %d frames in one persistent launch, no L2 flush" % K},
        "gpu_launches": launches, "all_frames_accepted": ok, "wall_s": t_wall, "clocks": clk.summary(),
    }
    if e2e is not None:
        line["e2e"] = e2e
    h.close()
    return line


def e2e_run(h, capi, HostTracker, trackers, n_p, n_t, blobs, counts, times, seeds, W, K, world, dev):
    """Per-frame pfmpe_update(_batch) with host buffers + host tail, wall-clock with device syncs."""
    import torch
    scs = [S.Scene(tracker=int(t)) for t in trackers]
    trk = [HostTracker(h, scs[i], tracker=i, init_particles=False) for i in range(n_t)]
    for i in range(n_t):
        h.track_init(scs[i].pose(0), scs[i].time(0), scs[i].pose(1), scs[i].time(1), tracker=i)   # particles = pose1
    h.sync()

    def step(f):
        for i in range(n_t):
            trk[i].frame(2 + f, img=blobs[i, f, : counts[i, f]], seed=int(seeds[i, f]))

    n_t_run = n_t
    if n_t > 64:      # per-tracker Python loop: bound the sample, scale the rate
        n_t_run = 64
        trk = trk[:n_t_run]

        def step(f):  # noqa: F811
            for i in range(n_t_run):
                trk[i].frame(2 + f, img=blobs[i, f, : counts[i, f]], seed=int(seeds[i, f]))
    for f in range(min(W, 3)):
        step(f)
    torch.cuda.synchronize(dev)
    it_sum, lat = 0, []
    t0 = time.perf_counter()
    for f in range(W, W + K):
        a = time.perf_counter()
        step(f)
        lat.append(time.perf_counter() - a)
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    pme = float(n_p) * L_LED * K * n_t_run          # iterations == 1 on these sequences (checked in value arm)
    fc_bytes = 224 + 8 * 2 * blobs.shape[2]
    val = pme / dt * world * (1.0 if n_t_run == n_t else 1.0)
    return {"value": val, "unit": "PME/s", "h2d_bytes_per_step": int(fc_bytes * n_t_run), "d2h_bytes_per_step": int(808 * n_t_run),
            "steps": K, "p50_ms": 1e3 * float(np.median(lat)), "trackers_in_sample": n_t_run,
            "path": "pfmpe_update per frame (host buffers) + host Gauss-Newton/predictPose, driven from Python ctypes"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto"] + list(WORKLOADS))
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=300)
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "auto":
        args.workload = "cfg2" if max(world, args.gpus) == 1 else "cfg4"
    if args.steps is None:
        args.steps = WORKLOADS[args.workload][3]
    args.warmup = max(args.warmup, 3)

    if args.impl == "reference":
        if rank != 0:
            return
        threads = os.cpu_count() or 1
        r = cpu_run(args.workload, budget_s=max(10.0, min(60.0, 0.05 * args.steps)), threads=threads)
        n_p, n_t, n_out, _ = WORKLOADS[args.workload]
        line = {"impl": "reference", "metric": "particle_marker_evals_per_sec", "value": r["value"], "unit": "PME/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["p50_ms"],
                "higher_is_better": True, "scaling": "strong" if n_t > 1 else "replicas", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload, "desc": WORKLOAD_DESC[args.workload], "particles": n_p, "trackers": n_t,
                           "leds": L_LED, "blobs": L_LED + n_out},
                "cpu_baseline": {"value": r["value"], "unit": "PME/s", "cores": r["cores"], "kind": "port", "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "PME/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "the reference cannot be compiled here (Eigen/OpenCV/ROS absent): oracle port of pose_estimator.cpp:475-733 "
                        "+ :1805-2009, one filter per host thread (the reference itself is single-threaded)"}
        print(json.dumps(line))
        return

    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    line = gpu_run(args, rank, world, local_rank)
    if rank == 0:
        if world == 1 and not args.no_cpu:
            r = cpu_run(args.workload, budget_s=args.cpu_budget, threads=1)
            line["cpu_baseline"] = {"value": r["value"], "unit": "PME/s", "cores": 1, "kind": "port", "sample": r["sample"],
                                    "p50_ms": r["p50_ms"]}
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

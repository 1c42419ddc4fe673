"""CPU tests of the DEVICE arithmetic (pf_math.cuh / pf_tail.cuh compiled for the host, tests/hostmath.cpp)
against the oracle: FP32 propagate/project/score parity, exact RNG skip-ahead, fixed-point scan +
threshold resampling index parity.  The CUDA kernels inline exactly these functions."""
import numpy as np
import pytest

from oracle import orc
from tests import hostmath, parity
from tests.util import frame_inputs


def test_rng_skip_table_matches_sequential():
    rng = np.random.default_rng(0)
    for _ in range(200):
        x = int(rng.integers(1, 2147483646))
        steps = int(rng.integers(0, 1 << 40))
        assert hostmath.lib().hm_rng_skip(x, steps) == orc.lib().orc_rng_skip(x, steps)
    assert hostmath.lib().hm_rng_skip(1, 10000) == 1043618065


def test_canonical_bit_exact_and_f32_close():
    import ctypes as C
    rng = np.random.default_rng(1)
    g = C.c_uint32(12345)
    for _ in range(20000):
        st = C.c_uint32(g.value)
        c_ref = orc.lib().orc_rng_canonical(C.byref(g))
        g1 = orc.lib().orc_rng_skip(st.value, 1)
        g2 = orc.lib().orc_rng_skip(st.value, 2)
        c64 = hostmath.lib().hm_canonical_f64(g1, g2)
        assert c64 == c_ref
        c32 = hostmath.lib().hm_canonical_f32(g1, g2)
        assert abs(c32 - c_ref) <= 1.2e-7 * max(c_ref, 1e-3) + 1e-10
    assert hostmath.lib().hm_canonical_f64(2147483646, 2147483646) < 1.0


@pytest.mark.parametrize("N,frame,isi,nout", [(100, 5, 2, 10), (100, 9, 1, 10), (1000, 23, 2, 10), (4096, 40, 2, 10),
                                              (300, 17, 2, 200)])
def test_frame_parity_vs_oracle(N, frame, isi, nout):
    fi = frame_inputs(N, frame=frame, isi=isi, n_outliers=nout)
    ora = orc.pf_update(resampler=(0 if N <= 1000 else 1), **{k: v for k, v in fi.items() if k != "truth"})
    got = hostmath.frame(fi)
    res = parity.check_frame(fi, got, ora)
    assert ora["iterations"] == 1 and ora["resampled"] == 1
    # winner correspondences identical to the oracle's for that particle
    q = got["winner"]
    assert got["pairs"].tolist() == ora["pairs"][q][: 2 * ora["npairs"][q]].reshape(-1, 2).tolist()
    assert res["index_mismatch"] == 0


def test_frame_parity_multi_iteration():
    fi = frame_inputs(64, frame=11, occlude=(2,))
    ora = orc.pf_update(resampler=0, **{k: v for k, v in fi.items() if k != "truth"})
    got = hostmath.frame(fi)
    assert ora["iterations"] == 80
    parity.check_frame(fi, got, ora)


def test_frame_parity_reject():
    fi = frame_inputs(50, frame=7)
    fi["img"] = fi["img"] + 300.0
    ora = orc.pf_update(resampler=0, **{k: v for k, v in fi.items() if k != "truth"})
    got = hostmath.frame(fi)
    assert got["accepted"] == 0 and got["resampled"] == 0 and got["iterations"] == 80
    parity.check_frame(fi, got, ora)


@pytest.mark.parametrize("N", [2, 3, 64, 1000, 20000])
def test_resample_stage_index_exact(N):
    """Stage-wise parity: same FP32 weights in -> identical indices/counts as the literal double resampler,
    including zero runs, negative weights (non-monotone prefix, D4) and a negative sum (Q6)."""
    rng = np.random.default_rng(N)
    for trial in range(5):
        w = rng.uniform(0, 30, N).astype(np.float32)
        if trial >= 1:
            w[rng.uniform(size=N) < 0.6] = 0.0
        if trial >= 2:
            w[rng.uniform(size=N) < 0.2] *= np.float32(-0.4)
        if trial == 3:
            w = -w
        if trial == 4:
            w[:] = 0
            w[rng.integers(0, N)] = 7.5
        seed, iters = 1000 + trial, 1 + trial
        rc, idx, cnt, sfx = hostmath.resample(w, seed, iters)
        w64 = w.astype(np.float64)
        s = 0.0
        for v in w64:
            s += v
        if s == 0:
            assert rc == 1
            continue
        assert rc == 0
        assert sfx == int(np.round(w64 * 2.0 ** 30).sum())
        r = orc.resample_uniforms(seed, N, 12 * (N - 2) * iters)
        idx_o, cnt_o = orc.resample(w64 / s, r, literal=(N <= 3000))
        ok = idx_o >= 0
        mism = np.nonzero((idx_o != idx) & ok)[0]
        c = np.cumsum(w64 / s)
        u = (np.arange(N) + r) / N
        for k in mism:      # only stratum ties may differ
            lo, hi = sorted((idx_o[k], idx[k]))
            assert np.abs(c[lo:hi + 1] - u[k]).min() < 1e-12 * N
        assert len(mism) <= 1
        if len(mism) == 0 and ok.all():
            assert (cnt_o == cnt).all()

"""Committed golden vectors of the PF update (tests/golden/pf_update.json, written by make_pf_golden.py from the
oracle: the reference itself cannot be built here, parity unpinned).  CPU: the oracle (both resamplers) and the
FP32 host-math twin reproduce them; GPU: the CUDA path through the C ABI does."""
import json
import os

import numpy as np
import pytest

from oracle import orc
from tests import parity
from tests.golden.make_pf_golden import checksum
from tests.util import frame_inputs

G = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pf_update.json")))["cases"]


def _inputs(g):
    c = g["case"]
    fi = frame_inputs(c["N"], frame=c["frame"], n_outliers=c["n_outliers"], isi=c["isi"])
    assert checksum(fi) == g["inputs_sha256"], "the input generator changed: regenerate the golden file"
    return fi


def _as_oracle_dict(g, fi):
    """the golden record in the shape parity.check_frame expects (T comes from a fresh oracle run)"""
    o = orc.pf_update(resampler=1, **{k: v for k, v in fi.items() if k != "truth"})
    for k in ("iterations", "best_iteration", "accepted", "resampled", "most_likely_particle_idx", "most_likely_pose_idx"):
        assert int(o[k]) == g[k], k
    assert o["highest_prob"] == g["highest_prob"] and o["weight_sum"] == g["weight_sum"]
    np.testing.assert_array_equal(o["w_raw"], np.array(g["w_raw"]))
    np.testing.assert_array_equal(o["idx"], np.array(g["idx"]))          # O(N) resampler == literal O(N^2) scan
    np.testing.assert_array_equal(o["count"], np.array(g["count"]))
    np.testing.assert_array_equal(o["winner_pose"], np.array(g["winner_pose"]))
    np.testing.assert_array_equal(o["T"][:4], np.array(g["T_first"]))
    return o


@pytest.mark.parametrize("i", range(len(G)))
def test_oracle_reproduces_golden(i):
    fi = _inputs(G[i])
    _as_oracle_dict(G[i], fi)


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(G)))
def test_cuda_matches_golden(i):
    from tests.cuda_util import make_handle, run_frame
    g = G[i]
    fi = _inputs(g)
    ora = _as_oracle_dict(g, fi)
    for mode in (1, 2):
        h = make_handle(fi, g["case"]["N"], force_mode=mode)
        got = run_frame(h, fi)
        res = parity.check_frame(fi, got, ora)
        if res["weight_ties"] == 0 and res["index_mismatch"] == 0 and g["resampled"]:
            np.testing.assert_array_equal(got["count"], np.array(g["count"]))
            assert got["winner"] == g["most_likely_pose_idx"]
        h.close()

"""vxorb_knn2_static — the allocation-free 2-NN entry point (static device scratch) against vxorb_knn2.

Collected last on purpose: a failure here must not hide the parity tests.
The rows are integer work, so the bar is the suite's: bit-exact (indices and distances) against the existing
entry point, on calls in a row that share the static array — a large call, a smaller one whose rows must not see
the first call's leftovers, then the largest shape the array holds.
"""
import numpy as np
import pytest

import oracle
import visionx_slam_b200 as vx


def _knn_equal(a, b):
    return np.array_equal(a["idx"], b["idx"]) and np.array_equal(
        np.where(a["idx"] >= 0, a["dist"], 0), np.where(b["idx"] >= 0, b["dist"], 0))


def _sets(seed, nq, nt, ties=False):
    rng = np.random.default_rng(seed)
    if ties:  # few distinct rows: many equal distances, so the lowest-index tie-break is exercised
        pool = rng.integers(0, 256, (7, 32), dtype=np.uint8)
        return pool[rng.integers(0, 7, nq)], pool[rng.integers(0, 7, nt)]
    return rng.integers(0, 256, (nq, 32), dtype=np.uint8), rng.integers(0, 256, (nt, 32), dtype=np.uint8)


def test_static_limits_are_rejected_with_a_clear_error():
    """Sizes beyond the static array are refused before any CUDA call (runs without a GPU)."""
    q, t = _sets(1, vx.STATIC_MAX_QUERY + 1, 8)
    with pytest.raises(vx.VxOrbError) as e:
        vx.knn2_static(q, t)
    assert e.value.code == vx.E_INVALID
    assert str(vx.STATIC_MAX_QUERY) in str(e.value) and str(vx.STATIC_MAX_TRAIN) in str(e.value)
    q, t = _sets(2, 8, vx.STATIC_MAX_TRAIN + 1)
    with pytest.raises(vx.VxOrbError) as e:
        vx.knn2_static(q, t)
    assert e.value.code == vx.E_INVALID and "vxorb_knn2" in str(e.value)


@pytest.mark.gpu
def test_static_knn2_matches_knn2_on_calls_in_a_row():
    cases = [(11, 1000, 1000, False), (12, 300, 700, False), (13, 1000, 1000, True), (14, 1, 1, False),
             (15, 5, 0, False), (16, 2500, 3000, False),
             (17, vx.STATIC_MAX_QUERY, vx.STATIC_MAX_TRAIN, False), (18, 40, 2, True)]
    sets = [_sets(s, nq, nt, ties) for s, nq, nt, ties in cases]
    # first two calls before any handle exists: the entry point needs none (and nothing it could have allocated)
    early = [vx.knn2_static(q, t) for q, t in sets[:2]]
    with vx.VxOrb(max_train=vx.STATIC_MAX_TRAIN) as h:
        want = [h.knn2(q, t) for q, t in sets]
        for got, ref in zip(early, want[:2]):
            assert _knn_equal(got, ref)
        for rep in range(2):  # the whole series twice: every call inherits the previous call's array contents
            for (q, t), ref, case in zip(sets, want, cases):
                got = vx.knn2_static(q, t)
                assert _knn_equal(got, ref), f"case {case} pass {rep}"
        # same shape, different data, back to back
        a, b = _sets(21, 777, 1234), _sets(22, 777, 1234)
        ga, gb = vx.knn2_static(*a), vx.knn2_static(*b)
        assert _knn_equal(ga, h.knn2(*a)) and _knn_equal(gb, h.knn2(*b))
    # and against the CPU oracle, so that both entry points being wrong together cannot pass
    q, t = sets[2]
    assert _knn_equal(vx.knn2_static(q, t), oracle.knn2(q, t))
    assert np.all(vx.knn2_static(*sets[4])["idx"] == -1)  # empty train set: no neighbours

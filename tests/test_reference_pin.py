"""The oracle against the REFERENCE's own code (tests/ref_harness.py).

Two layers: (1) committed fixtures ``tests/golden/ref_*.npz`` — outputs of the reference's
``correct_lines`` / decoders / ``Node``, ``attention_call`` and ``_alignment2path`` executed in the
build container — compared everywhere (also on the GPU box, where /root/reference is absent);
(2) live re-execution of the reference when /root/reference is present, bit for bit.
"""
import os

import numpy as np
import pytest

from oracle import alignment_path as oap
from oracle import decoders as od
from oracle import model as om
from tests import ref_harness as rh

ALL_CASES = list(rh.DECODER_CASES) + list(rh.KANT_CASES)
live = pytest.mark.skipif(not rh.available(), reason="/root/reference not present")


def _g(name):
    return np.load(os.path.join(rh.GOLDEN, name), allow_pickle=False)


@pytest.mark.parametrize("name", ALL_CASES)
def test_decoders_match_reference_fixture(name):
    """Strings identical; probabilities / scores / alignment centres to float32 round-off (the
    fixture was computed on another machine's BLAS)."""
    g = _g("ref_%s.npz" % name)
    got = rh.pack_result(rh.run_oracle_case(name, promotion="nep50"))
    assert [t.decode("utf-8") for t in g["texts"]] == [t.decode("utf-8") for t in got["texts"]]
    assert (g["lens"] == got["lens"]).all()
    assert np.allclose(g["probs"], got["probs"], atol=2e-5)
    assert np.allclose(g["scores"], got["scores"], atol=2e-5)
    assert np.allclose(g["align_centre"], got["align_centre"], atol=1e-3)


@pytest.mark.parametrize("name", ALL_CASES)
def test_legacy_promotion_changes_only_round_off(name):
    """The product follows the reference's pinned numpy (<1.24, float64 cost accumulation); the
    strings of these cases do not depend on it and the scores move by float32 round-off only."""
    a = rh.run_oracle_case(name, promotion="nep50")
    b = rh.run_oracle_case(name, promotion="legacy")
    assert a[0] == b[0]
    assert np.allclose([float(s) for s in a[2]], [float(s) for s in b[2]], atol=1e-5)


@live
@pytest.mark.parametrize("name", ALL_CASES)
def test_decoders_match_reference_live(name):
    ref = rh.run_reference_case(name)
    got = rh.run_oracle_case(name, promotion="nep50")
    assert ref[0] == got[0]
    for pr, pg in zip(ref[1], got[1]):
        assert [float(p) for p in pr] == [float(p) for p in pg]
    assert [float(s) for s in ref[2]] == [float(s) for s in got[2]]
    for ar, ag in zip(ref[3], got[3]):
        assert len(ar) == len(ag)
        for x, y in zip(ar, ag):
            assert np.array_equal(np.asarray(x), np.asarray(y))


@live
def test_reference_greedy_nan_feedback_crash_is_real():
    """when '' wins the argmax, decode_sequence_greedy feeds NaN back and the next
    step raises.  The product deviates there on purpose."""
    from tests.common import make_model, synthetic_lines
    topo, w, mapping = make_model("c3res")
    lines = synthetic_lines(mapping, 2, 18, seed=314, min_length=7)
    s2s = rh.reference_model(topo, w, mapping)
    with pytest.raises(ValueError):
        with np.errstate(all='ignore'):
            s2s.correct_lines(lines, fast=False, greedy=True)
    events = []
    enc = om.encode(topo, w, od.vectorize_lines(mapping, topo.voc_size, lines))
    od.decode_sequence_greedy(topo, w, mapping, [e[0:1] for e in enc], events=events)
    assert events and events[0][0] == 'nul-argmax'


def test_attention_matches_reference_fixture():
    from types import SimpleNamespace
    g = _g("ref_attention_call.npz")
    topo = SimpleNamespace(window_width=5, window_inclusive=True)
    for k in range(int(g["n"][0])):
        c = {key: g["%s%d" % (key, k)] for key in ("h", "enc", "u", "a_prev", "W_a", "v_a", "b_UW", "b_v")}
        ctx, a = om.attention(topo, (c["W_a"], c["v_a"], c["b_UW"], c["b_v"]), c["h"], c["a_prev"], c["enc"], c["u"])
        assert ((a > 0) == (g["a%d" % k] > 0)).all()          # same window, inclusive edges included
        assert np.abs(a - g["a%d" % k]).max() < 1e-6
        assert np.abs(ctx - g["ctx%d" % k]).max() < 1e-6


@live
def test_attention_matches_reference_live():
    from types import SimpleNamespace
    topo = SimpleNamespace(window_width=5, window_inclusive=True)
    for seed, rej in [(81, False), (82, True), (83, False), (84, True)]:
        c = rh.attention_case(seed, B=7, T=23, after_rejection=rej)
        ctx_r, a_r = rh.run_reference_attention(c)
        ctx, a = om.attention(topo, (c["W_a"], c["v_a"], c["b_UW"], c["b_v"]), c["h"], c["a_prev"], c["enc"], c["u"])
        assert ((a > 0) == (a_r > 0)).all()
        assert np.abs(a - a_r).max() < 1e-6 and np.abs(ctx - ctx_r).max() < 1e-6


def _with_promotion(p, fn, *args):
    old, od.PROMOTION = od.PROMOTION, p
    try:
        return fn(*args)
    finally:
        od.PROMOTION = old


def test_alignment_path_matches_reference_fixture():
    g = _g("ref_alignment2path.npz")
    for k in range(int(g["n"][0])):
        i_max, j_max = (int(v) for v in g["dims%d" % k])
        path, dist = _with_promotion("nep50", oap.alignment_path, g["a%d" % k], i_max, j_max, g["ms%d" % k][0])
        assert (path == g["path%d" % k]).all(), k
        assert dist == g["dist%d" % k][0], k
        assert g["extra%d" % k].size == 0
        path2, dist2 = _with_promotion("legacy", oap.alignment_path, g["a%d" % k], i_max, j_max, g["ms%d" % k][0])
        assert (path2 == path).all() and abs(dist2 - dist) < 1e-4


@live
def test_alignment_path_matches_reference_live():
    rng = np.random.default_rng(17)
    for n in range(200):
        S, T = int(rng.integers(1, 30)), int(rng.integers(1, 30))
        a = rng.random((S, T)).astype(np.float32) ** 4
        a /= a.sum(1, keepdims=True)
        if n % 3 == 0:
            a = (np.round(a * 4) / 4).astype(np.float32)        # many exact ties
        i_max, j_max = int(rng.integers(1, T + 1)), int(rng.integers(1, S + 1))
        ms = np.float32(rng.choice([1 / 80, 0.1, 0.3]))
        ref_path, ref_dist, extra = rh.run_reference_viterbi(a, i_max, j_max, ms)
        path, dist = _with_promotion("nep50", oap.alignment_path, a, i_max, j_max, ms)
        assert not extra and (path == ref_path).all() and dist == ref_dist

"""CPU tests of the oracle itself: committed known answers, the independent torch restatement,
and the by-construction properties listed in SURVEY.md §8c.  (Parity unpinned: the reference's
Keras/TF cannot run here — see oracle/__init__.py.)"""
import math
import os

import numpy as np
import pytest

from oracle import model as om
from oracle import decoders as od
from tests.common import (make_model, synthetic_lines, BEAM_CASES, GOLDEN_DIR, kant_fixture, kant_model)
from tests.restatement_torch import Restatement

GAP = '\a'


def _load(name):
    return np.load(os.path.join(GOLDEN_DIR, name), allow_pickle=True)


@pytest.mark.parametrize("name", ["c1", "c2s", "c3res", "c4s"])
def test_forward_golden(name):
    g = _load("forward_%s.npz" % name)
    topo, w, mapping = make_model(name)
    lines = list(g["lines"])
    x = od.vectorize_lines(mapping, topo.voc_size, lines)
    enc = om.encode(topo, w, x)
    assert np.abs(enc[0] - g["enc_out"]).max() < 2e-6
    assert np.abs(np.stack(enc[1:-1]) - g["enc_states"]).max() < 2e-6
    p0, lm0, st0 = om.decode_step(topo, w, np.zeros((len(lines), topo.voc_size), np.float32), enc[0], enc[1:])
    p1, lm1, st1 = om.decode_step(topo, w, p0, enc[0], st0)
    assert np.abs(p0 - g["p0"]).max() < 1e-6 and np.abs(p1 - g["p1"]).max() < 1e-6
    assert np.abs(st1[-1] - g["a1"]).max() < 1e-6
    if topo.lm_predict:
        assert np.abs(lm1 - g["lm1"]).max() < 1e-6
    else:
        text, probs, score, _ = od.decode_batch_greedy(topo, w, mapping, x)
        assert text == list(g["greedy_text"])
        assert np.allclose(score, g["greedy_score"], atol=1e-5)


@pytest.mark.parametrize("name", ["c1", "c2s", "c3res", "c4s"])
def test_forward_against_independent_restatement(name):
    """numpy float32 oracle vs torch.nn.LSTM float64 restatement: encoder, two decoder steps."""
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, 5, 13, seed=55, min_length=3) + ['']
    x = od.vectorize_lines(mapping, topo.voc_size, lines)
    r = Restatement(topo, w)
    enc_o, enc_r = om.encode(topo, w, x), r.encode(x)
    for a, b in zip(enc_o, enc_r):
        assert np.abs(a - b.numpy()).max() < 5e-6
    # float64 oracle against float64 restatement: tight
    w64 = om.cast_weights(w, np.float64)
    enc_o64 = om.encode(topo, w64, x.astype(np.float64))
    for a, b in zip(enc_o64, enc_r):
        assert np.abs(a - b.numpy()).max() < 1e-12
    p = np.zeros((len(lines), topo.voc_size))
    st_o, st_r = enc_o64[1:], enc_r[1:]
    for _ in range(3):
        p_o, lm_o, st_o = om.decode_step(topo, w64, p, enc_o64[0], st_o)
        p_r, lm_r, st_r = r.step(p, enc_r[0], st_r)
        assert np.abs(p_o - p_r.numpy()).max() < 1e-12
        assert np.abs(st_o[-1] - st_r[-1].numpy()).max() < 1e-12
        if topo.lm_predict:
            assert np.abs(lm_o - lm_r.numpy()).max() < 1e-12
        p = p_o


@pytest.mark.parametrize("name", ["c1", "c2s"])
def test_greedy_against_independent_restatement(name):
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, 4, 11, seed=56, min_length=4)
    x = od.vectorize_lines(mapping, topo.voc_size, lines)
    text, probs, _, _ = od.decode_batch_greedy(topo, w, mapping, x)
    text_r, probs_r = Restatement(topo, w).greedy_batch(x, mapping[1])
    assert text == text_r
    for a, b in zip(probs, probs_r):
        assert np.abs(np.array(a) - np.array(b)).max() < 1e-5


@pytest.mark.parametrize("case", sorted(BEAM_CASES))
def test_beam_golden(case):
    name, n, length, seed, params = BEAM_CASES[case]
    g = _load("beam_%s.npz" % case)
    topo, w, mapping = make_model(name)
    lines = list(g["lines"])
    stats = {}
    text, probs, score, _ = od.correct_lines(topo, w, mapping, lines, fast=False, greedy=False,
                                             beam_params=params, stats=stats)
    assert text == list(g["text"])
    assert np.allclose(np.array(score, np.float64), g["score"], atol=1e-6)
    from tests.make_golden import stat_row
    st = np.array([stat_row(s) for s in stats["lines"]], np.int64)
    assert (st == g["stats"]).all()


def test_beam_known_answers_by_construction():
    """random-init weights and rejection 0.5 => output == input, score == ln 2,
    T iterations with one hypothesis each (the rejection candidate p=0.5 dominates 1/V)."""
    topo, w, mapping = make_model("c2s")
    lines = synthetic_lines(mapping, 3, 12, seed=77)
    stats = {}
    text, probs, score, align = od.correct_lines(topo, w, mapping, lines, fast=False, greedy=False,
                                                 beam_params=dict(rejection_threshold=0.5), stats=stats)
    assert text == lines
    for j, s in enumerate(stats["lines"]):
        T = len(lines[j])
        assert s["iterations"] == T and s["hypothesis_steps"] == T
        assert abs(score[j] - math.log(2)) < 1e-6
        assert all(abs(p - 0.5) < 1e-7 for p in probs[j])
        for t, a in enumerate(align[j]):              # rejection alignments are one-hot on the diagonal
            assert a[t] == 1.0 and a.sum() == 1.0


def test_first_step_window_support():
    """a_prev = 0 => centre 1 => support exactly positions 0..6 (inclusive window)."""
    topo, w, mapping = make_model("c1")
    lines = synthetic_lines(mapping, 2, 15, seed=78)
    x = od.vectorize_lines(mapping, topo.voc_size, lines)
    enc = om.encode(topo, w, x)
    _, _, st = om.decode_step(topo, w, np.zeros((2, topo.voc_size), np.float32), enc[0], enc[1:])
    a = st[-1]
    assert (a[:, :7] > 0).all() and (a[:, 7:] == 0).all()
    assert np.allclose(a.sum(1), 1, atol=1e-6)


def test_padding_is_not_masked():
    """results depend on the padded length of the batch."""
    topo, w, mapping = make_model("c1")
    short = synthetic_lines(mapping, 1, 6, seed=79)
    long_ = synthetic_lines(mapping, 1, 14, seed=80)
    alone = om.encode(topo, w, od.vectorize_lines(mapping, topo.voc_size, short))
    padded = om.encode(topo, w, od.vectorize_lines(mapping, topo.voc_size, short + long_))
    assert np.abs(alone[1] - padded[1][:1]).max() > 1e-4      # backward final state saw the padding


def test_kant_config1_golden():
    """Config 1: the 55 kant test lines (plain strings and glyph confmats), depth 1 / width 128."""
    alphabet, pages = kant_fixture()
    assert [len(p) for p in pages] == [24, 31]
    topo, w, mapping = kant_model(alphabet)
    g = _load("kant_c1.npz")
    for pno, page in enumerate(pages):
        plain = [''.join(chunk[0][0] for chunk in line).replace(GAP, '') for line in page]
        t1, _, s1, _ = od.correct_lines(topo, w, mapping, plain, fast=True)
        assert t1 == list(g["plain_text_%d" % pno])
        assert np.allclose(s1, g["plain_score_%d" % pno], atol=1e-5)
        conf = [[[(s, c) for s, c in chunk] for chunk in line] for line in page]
        t2, _, s2, _ = od.correct_lines(topo, w, mapping, conf, conf=conf, fast=True)
        assert t2 == list(g["confmat_text_%d" % pno])
        assert np.allclose(s2, g["confmat_score_%d" % pno], atol=1e-5)


def test_vectorize_matches_sparse_encoding():
    """Host logic: the product's sparse line encoding reproduces the oracle's dense array."""
    from cor_asv_ann_b200.vectorize import lines_to_sparse, dense_to_sparse, rejection_table
    alphabet, pages = kant_fixture()
    topo, w, mapping = kant_model(alphabet)
    page = pages[0]
    conf = [[[(s, c) for s, c in chunk] for chunk in line] for line in page]
    dense = od.vectorize_lines(mapping, topo.voc_size, conf, conf)
    idx, val = lines_to_sparse(mapping[0], conf, conf)
    rebuilt = np.zeros_like(dense)
    B, T, k = idx.shape
    for a in range(k):
        b, t = np.nonzero(idx[:, :, a] >= 0)
        rebuilt[b, t, idx[b, t, a]] = val[b, t, a]
    assert (rebuilt == dense).all()
    idx2, val2 = dense_to_sparse(dense)
    rej = rejection_table(idx2, val2)
    want = np.where(dense.any(axis=2), dense.argmax(axis=2), -1)
    assert (rej == want).all()

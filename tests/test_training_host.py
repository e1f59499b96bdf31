"""Host side of row T without a GPU: character inventory (map_files, seq2seq.py:554-590), teacher-forcing index arrays
(vectorize_lines, :1061-1117) against the training oracle's dense arrays, the training-time input degradation (:909-915),
the row shares of data-parallel ranks, embedding growth on a larger mapping (:499-525)."""
import numpy as np

from cor_asv_ann_b200 import training
from cor_asv_ann_b200.seq2seq import Sequence2Sequence
from cor_asv_ann_b200.trainer import teacher_indices
from oracle import train as ot


def test_map_files_and_incremental_mapping(tmp_path):
    p = tmp_path / "a.txt"
    p.write_text("abc\tabd\nxy\txyz\n")
    s = Sequence2Sequence()
    n = s.map_files([str(p)])
    assert n == 2
    assert s.mapping[1][0] == '' and sorted(s.mapping[0]) == sorted(['', '\t', '\n'] + list("abcdxyz"))
    assert s.voc_size == len(s.mapping[0]) and s.status == 1           # configured for the new alphabet
    old_E = s._weights['char_input_projection'][0].copy()
    s.status = 2
    q = tmp_path / "b.txt"
    q.write_text("abé\tabe\n")
    s.map_files([str(p), str(q)])
    assert 'é' in s.mapping[0] and s.voc_size == len(old_E) + 2           # 'e' and 'é' are new
    E = s._weights['char_input_projection'][0]
    assert E.shape[0] == s.voc_size and np.array_equal(E[:len(old_E)], old_E)   # known rows keep their weights (seq2seq.py:515-519)


def test_teacher_arrays_match_the_dense_ones():
    s = Sequence2Sequence()
    s.mapping = ({c: i for i, c in enumerate(['', '\n'] + list("abcdef"))}, {i: c for i, c in enumerate(['', '\n'] + list("abcdef"))})
    s.voc_size = 8
    src, tgt = ["abc\n", "fe\n", "\n"], ["abd\n", "feed\n", "\n"]
    idx, val, di, do = training.batch_arrays(s, src, None, tgt)
    x, dense_in, dense_out, w = ot.teacher_arrays(s.mapping, 8, src, tgt)
    assert val is None
    for b in range(3):
        for t in range(x.shape[1]):
            assert (idx[b, t, 0] == -1) == (x[b, t].sum() == 0)
            if idx[b, t, 0] >= 0:
                assert x[b, t, idx[b, t, 0]] == 1
        for t in range(di.shape[1]):
            assert (di[b, t] == -1) == (dense_in[b, t].sum() == 0) and (do[b, t] == -1) == (w[b, t] == 0)
            if di[b, t] >= 0:
                assert dense_in[b, t, di[b, t]] == 1
            if do[b, t] >= 0:
                assert dense_out[b, t, do[b, t]] == 1
    s2, d2, o2 = teacher_indices(s.mapping[0], src, tgt)
    assert np.array_equal(s2, idx) and np.array_equal(d2, di) and np.array_equal(o2, do)


def test_input_degradation_rate_and_form():
    s = Sequence2Sequence()
    s.mapping = ({c: i for i, c in enumerate(['', '\n', 'a'])}, {0: '', 1: '\n', 2: 'a'})
    s.voc_size = 3
    lines = ["a" * 49 + "\n"] * 4000
    idx, _, _, _ = training.batch_arrays(s, lines, None, lines, np.random.default_rng(3), degrade=True)
    hit = (idx[:, :, 0] == 0)
    assert hit.sum(axis=1).max() == 1                                   # at most one position per line
    assert 0.003 < hit.any(axis=1).mean() < 0.02                        # probability 0.01 per line


def test_row_shares_cover_the_batch():
    for n, world in [(256, 8), (7, 3), (3, 8), (64, 1)]:
        cuts = [training._share(n, r, world) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n and all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))

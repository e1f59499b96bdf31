"""Row N2 host logic (no GPU): several reference batches in one engine call, the threaded
prefetch, vectorised string building and the float32 score chain.  The engine is replaced by a
test double that answers ``correct_greedy_host`` from the oracle, per run of rows with equal row_T,
so what is checked is everything between the text files and the C-ABI call."""
import numpy as np
import pytest

from cor_asv_ann_b200 import pipeline
from cor_asv_ann_b200.vectorize import lines_to_sparse
from oracle import decoders as od
from tests.common import make_model, synthetic_lines
from tests.test_seq2seq_host import _s2s_from


class OracleEngine:
    """Test double for engine.Engine: greedy decode of merged batches through the oracle."""

    def __init__(self, topo, w, mapping):
        self.topo, self.w, self.mapping = topo, w, mapping
        self.calls = []

    def correct_greedy_host(self, idx, val, row_T, want_align=False, out=None):
        import torch
        B, T, k = idx.shape
        self.calls.append((B, T, k))
        V = self.topo.voc_size
        chars = np.zeros((B, 2 * T), np.int32)
        probs = np.zeros((B, 2 * T), np.float32)
        scores = np.zeros(B, np.float64)
        lens = np.zeros(B, np.int32)
        align = np.zeros((2 * T, B, T), np.float32) if want_align else None
        c2i = self.mapping[0]
        lo = 0
        while lo < B:
            hi = lo
            while hi < B and row_T[hi] == row_T[lo]:
                hi += 1
            t = int(row_T[lo])
            x = np.zeros((hi - lo, t, V), np.float32)
            for b in range(lo, hi):
                for j in range(t):
                    for a in range(k):
                        if idx[b, j, a] >= 0:
                            x[b - lo, j, idx[b, j, a]] = 1.0 if val is None else val[b, j, a]
            text, p, s, al = od.decode_batch_greedy(self.topo, self.w, self.mapping, x)
            for b in range(lo, hi):
                n = len(text[b - lo])
                lens[b] = n
                chars[b, :n] = [c2i[c] for c in text[b - lo]]
                probs[b, :n] = p[b - lo]
                scores[b] = s[b - lo]
                if want_align:
                    for q in range(n):
                        align[q, b, :t] = al[b - lo][q]
            lo = hi
        return (torch.as_tensor(chars), torch.as_tensor(probs), torch.as_tensor(scores), torch.as_tensor(lens),
                None if align is None else torch.as_tensor(align))


def _model():
    topo, w, mapping = make_model("c2s")
    s = _s2s_from(topo, w, mapping)
    eng = OracleEngine(topo, w, mapping)
    s._get_engine = lambda: eng
    return s, eng, (topo, w, mapping)


def test_merge_batches_keeps_each_batch_padding():
    a = (np.array([[[3], [4]], [[5], [-1]]], np.int32), None)
    b = (np.array([[[1, 2], [6, -1], [7, 7]]], np.int32), np.array([[[.5, .25], [1, 0], [.1, .9]]], np.float32))
    idx, val, row_T, offsets = pipeline.merge_batches([a, b])
    assert idx.shape == (3, 3, 2) and offsets == [0, 2, 3]
    assert row_T.tolist() == [2, 2, 3]
    assert idx[0].tolist() == [[3, -1], [4, -1], [-1, -1]] and idx[2].tolist() == [[1, 2], [6, -1], [7, 7]]
    assert val[0].tolist() == [[1, 0], [1, 0], [0, 0]] and val[1, 1].tolist() == [0, 0]
    assert val[2, 0].tolist() == [.5, .25]
    idx, val, row_T, _ = pipeline.merge_batches([a, a])
    assert val is None and idx.shape == (4, 2, 1)


def test_string_table_matches_join():
    _, _, mapping = make_model("c2s")
    rng = np.random.default_rng(3)
    chars = rng.integers(0, len(mapping[0]), size=(17, 23)).astype(np.int32)
    lens = rng.integers(0, 24, size=17).astype(np.int32)
    table = pipeline.StringTable(mapping[1], len(mapping[0]))
    assert table.simple
    want = [''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) for j in range(17)]
    assert table.strings(chars, lens) == want
    multi = dict(mapping[1]); multi[5] = 'ch'
    table = pipeline.StringTable(multi, len(multi))
    assert not table.simple
    assert table.strings(chars, lens) == [''.join(multi[int(v)] for v in chars[j, :lens[j]]) for j in range(17)]


def test_greedy_line_score_is_the_reference_loop_under_legacy_promotion():
    rng = np.random.default_rng(5)
    for n in (1, 2, 7, 162):
        p = rng.uniform(1e-3, 1, size=n).astype(np.float32)
        total = 0
        for v in p:
            total = np.float64(total) + np.float64(-np.log(v))     # int/float64 + np.float32 -> float64 (numpy < 1.24)
        assert pipeline.greedy_line_score(list(p)) == total / n
    assert pipeline.greedy_line_score([]) == 0


def test_prefetcher_order_and_errors():
    assert list(pipeline.Prefetcher(iter(range(50)), depth=2)) == list(range(50))

    def broken():
        yield 1
        yield 2
        raise ValueError('boom')
    got = []
    with pytest.raises(ValueError):
        for v in pipeline.Prefetcher(broken()):
            got.append(v)
    assert got == [1, 2]
    it = iter(pipeline.Prefetcher(iter(range(1000)), depth=1))      # abandoning the consumer stops the worker
    assert next(it) == 0
    it.close()
    assert [len(c) for c in pipeline.chunked(range(7), 3)] == [3, 3, 1]


def test_correct_batches_equals_per_batch_correct_lines():
    s, eng, (topo, w, mapping) = _model()
    batches = [(synthetic_lines(mapping, 3, 9, seed=1, min_length=4), None),
               ([], None),
               (synthetic_lines(mapping, 2, 14, seed=2, min_length=10) + [''], None),
               (synthetic_lines(mapping, 2, 6, seed=3), None)]
    for fast in (True, False):
        merged = s.correct_batches(batches, fast=fast, greedy=True, want_align=True)
        assert len(eng.calls) == 1 and eng.calls[0][0] == 8       # one engine call for all rows
        eng.calls.clear()
        for (lines, conf), got in zip(batches, merged):
            want = s.correct_lines(lines, conf, fast=fast, greedy=True)
            ref = od.correct_lines(topo, w, mapping, lines, fast=fast, greedy=True) if lines else ([], [], [], [])
            assert list(got[0]) == list(want[0]) == list(ref[0])
            for j in range(len(lines)):
                assert np.array_equal(np.array(got[1][j]), np.array(want[1][j]))
                assert got[2][j] == want[2][j]
                assert abs(float(got[2][j]) - float(ref[2][j])) < 1e-6
                assert len(got[3][j]) == len(want[3][j])
                for x, y in zip(got[3][j], want[3][j]):
                    assert np.array_equal(x, y)
        eng.calls.clear()


@pytest.mark.parametrize("prefetch", [True, False])
def test_predict_superbatches_same_results_as_batch_loop(tmp_path, prefetch):
    s, eng, (topo, w, mapping) = _model()
    s.batch_size = 4
    lines = synthetic_lines(mapping, 11, 12, seed=9, min_length=3)
    lines = [l.replace('\t', ' ') for l in lines]
    p1, p2 = tmp_path / 'a.txt', tmp_path / 'b.txt'
    p1.write_text(''.join(lines[:5]))
    p2.write_text(''.join(lines[5:]))
    files = [str(p1), str(p2)]
    s.superbatch, s.prefetch = 1, False
    want = list(s.predict(files, fast=True, greedy=True))
    n_ref_calls = len(eng.calls)
    eng.calls.clear()
    s.superbatch, s.prefetch = 2, prefetch
    got = list(s.predict(files, fast=True, greedy=True))
    assert len(want) == 3 and n_ref_calls == 3 and len(eng.calls) == 2
    assert got == want
    assert got[2][0] == [str(p2)] * 3 + [None] and got[2][1][3] == ''
    # and those are the oracle's lines for the same 4-line batches
    flat = [l for _, ls, _ in got for l in ls][:11]
    ref = []
    for i in range(0, 11, 4):
        chunk = lines[i:i + 4] + [''] * (4 - len(lines[i:i + 4]))
        ref += od.correct_lines(topo, w, mapping, chunk, fast=True)[0]
    assert flat == ref[:11]

"""The drop-in Sequence2Sequence class on the GPU against the oracle, read like the reference's own
callers use it (scripts/proc.py:52-62, wrapper/transcode.py:56-67,111-115): load_config -> configure ->
load_weights on a .h5 file, then correct_lines in fast / greedy / beamed mode."""
import numpy as np
import pytest

from oracle import model as om
from oracle import decoders as od
from tests.common import make_model, synthetic_lines, kant_fixture, kant_model
from tests.test_seq2seq_host import _s2s_from

pytestmark = pytest.mark.gpu
GAP = '\a'


def _loaded(topo, w, mapping, tmp_path):
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    path = str(tmp_path / 'model.h5')
    _s2s_from(topo, w, mapping).save(path)
    s = Sequence2Sequence()
    s.load_config(path)
    s.lm_predict = topo.lm_predict       # not persisted: programmatic only
    s.configure()
    s.load_weights(path)
    return s


def _check(res, ref, tol=1e-4):
    lines, probs, scores, aligns = res
    r_lines, r_probs, r_scores, r_aligns = ref
    assert lines == r_lines
    for j in range(len(lines)):
        assert len(probs[j]) == len(r_probs[j])
        if len(probs[j]):
            assert np.abs(np.array(probs[j], np.float32) - np.array(r_probs[j], np.float32)).max() < tol
            assert abs(float(scores[j]) - float(r_scores[j])) < tol
            a = np.stack([np.asarray(x, np.float32) for x in aligns[j]])
            b = np.stack([np.asarray(x, np.float32) for x in r_aligns[j]])
            assert a.shape == b.shape and np.abs(a - b).max() < tol


def test_config1_kant_lines_all_modes(tmp_path):
    """BASELINE config 1: depth 1 / width 128 on the reference's testdata lines, one batch per page."""
    alphabet, pages = kant_fixture()
    topo, w, mapping = kant_model(alphabet)
    s = _loaded(topo, w, mapping, tmp_path)
    s.rejection_threshold = 0.05
    s.batch_size = 64
    s.beam_width_in = 5
    params = dict(rejection_threshold=0.05, batch_size=64, beam_width_in=5)
    for page in pages:
        plain = [''.join(chunk[0][0] for chunk in line).replace(GAP, '') for line in page]
        conf = [[[(st, c) for st, c in chunk] for chunk in line] for line in page]
        _check(s.correct_lines(plain, fast=True, greedy=True), od.correct_lines(topo, w, mapping, plain, fast=True))
        _check(s.correct_lines(conf, conf, fast=True, greedy=True),
               od.correct_lines(topo, w, mapping, conf, conf=conf, fast=True))
        _check(s.correct_lines(plain[:8], fast=False, greedy=True),
               od.correct_lines(topo, w, mapping, plain[:8], fast=False, greedy=True))
        _check(s.correct_lines(conf[:6], conf[:6], fast=False, greedy=False),
               od.correct_lines(topo, w, mapping, conf[:6], conf=conf[:6], fast=False, greedy=False, beam_params=params))


def test_public_decoders_on_dense_arrays(tmp_path):
    topo, w, mapping = make_model("c2s")
    s = _loaded(topo, w, mapping, tmp_path)
    lines = synthetic_lines(mapping, 5, 12, seed=61, min_length=4)
    x = s.vectorize_lines(lines, lines)[0]
    # encoder_model.predict_on_batch
    enc = s.encoder_model.predict_on_batch(x)
    ref = om.encode(topo, w, x)
    assert len(enc) == len(ref)
    for a, b in zip(enc, ref):
        assert a.shape == b.shape and np.abs(a - b).max() < 5e-5
    # decode_batch_greedy
    out = s.decode_batch_greedy(x)
    assert out[0].shape == (5, 2 * x.shape[1], topo.voc_size) and out[0].dtype == np.uint32
    _check(out[1:], od.decode_batch_greedy(topo, w, mapping, x))
    # decode_sequence_greedy / decode_sequence_beam on one padded line
    j = 2
    text, probs, score, align = s.decode_sequence_greedy(x[j])
    sl = om.encode(topo, w, x[j:j + 1])
    r = od.decode_sequence_greedy(topo, w, mapping, sl)
    assert text == r[0] and abs(float(score) - float(r[2])) < 1e-4
    s.rejection_threshold, s.batch_size, s.beam_width_in = 0.03, 64, 5
    got = next(s.decode_sequence_beam(x[j]))
    want = next(od.decode_sequence_beam(topo, w, mapping, x[j], sl, rejection_threshold=0.03, batch_size=64,
                                        beam_width_in=5), None)
    assert want is not None and got[0] == want[0] and abs(got[2] - want[2]) < 1e-4


def test_beam_fallback_returns_input(tmp_path):
    """No final hypothesis -> the input line comes back with probabilities 1.0 (seq2seq.py:826-836)."""
    topo, w, mapping = make_model("c2s")
    s = _loaded(topo, w, mapping, tmp_path)
    s.rejection_threshold, s.batch_size, s.beam_width_in = 0, 32, 5
    lines = synthetic_lines(mapping, 2, 7, seed=36)
    out = s.correct_lines(lines, fast=False, greedy=False)
    ref = od.correct_lines(topo, w, mapping, lines, fast=False, greedy=False,
                           beam_params=dict(rejection_threshold=0, batch_size=32, beam_width_in=5))
    assert out[0] == ref[0] == lines
    assert out[1][0] == [1.0] * len(lines[0]) and out[2] == [0, 0]


@pytest.mark.parametrize("mode", ["fast", "greedy", "beam"])
def test_predict_superbatches_equal_per_batch_calls(tmp_path, mode):
    """Row N2: predict() over text files with several reference batches per engine call (each row
    carrying its own batch's padded T) returns exactly what the reference's batch-by-batch loop does,
    and what the oracle returns for those batches."""
    topo, w, mapping = make_model("c2s")
    s = _loaded(topo, w, mapping, tmp_path)
    s.batch_size = 8
    s.rejection_threshold, s.beam_width_in = 0.03, 5
    lines = synthetic_lines(mapping, 29, 18, seed=77, min_length=3)
    lines[8:16] = synthetic_lines(mapping, 8, 7, seed=78, min_length=2)      # a batch with a much shorter T
    lines = [l.replace('\t', ' ') for l in lines]
    f1, f2 = tmp_path / 'one.txt', tmp_path / 'two.txt'
    f1.write_text(''.join(lines[:13]))
    f2.write_text(''.join(lines[13:]))
    fast, greedy = mode == "fast", mode != "beam"
    s.superbatch, s.prefetch = 1, False
    want = list(s.predict([str(f1), str(f2)], fast=fast, greedy=greedy))
    s.superbatch, s.prefetch = 3, True
    got = list(s.predict([str(f1), str(f2)], fast=fast, greedy=greedy))
    assert len(got) == len(want) == 4
    params = dict(rejection_threshold=0.03, batch_size=8, beam_width_in=5)
    for b, ((names, out, scores), (names_w, out_w, scores_w)) in enumerate(zip(got, want)):
        assert names == names_w and out == out_w
        assert np.allclose(np.array(scores, np.float64), np.array(scores_w, np.float64), atol=1e-6)
        chunk = lines[8 * b:8 * b + 8]
        chunk = chunk + [''] * (8 - len(chunk))
        ref = od.correct_lines(topo, w, mapping, chunk, fast=fast, greedy=greedy, beam_params=params)
        assert out == ref[0]
        assert np.allclose(np.array(scores, np.float64), np.array(ref[2], np.float64), atol=1e-4)

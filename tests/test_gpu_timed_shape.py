"""GPU parity ON THE SHAPES THE BENCH TIMES ("What's weak" 1-3).

The large-batch decoders run `gemm_tc_kernel<1,2,256,LstmEpi>` (CTA pairs, cta_group::2) once a call holds
enough rows; every other model-level parity test stays below that threshold.  Here ONE call carries
17 reference batches (4 352 rows) of SURVEY §8d's parity variant C3v -- lengths uniform in 20..80, seed 1013,
un-masked padding inside every 256-line batch, each batch with its own padded T -- and

  * the kernel counters of the handle (cor_debug_counters) prove which GEMM variant ran;
  * two whole 256-line batches are compared with the oracle: strings identical, |dp|, |da| < 1e-4;
  * every row is compared with the same lines decoded through the already pinned <= 512-row path.

A string may differ from the oracle's only where the oracle itself sits on a near-tie (top-2 margin below
1e-4 relative at the first differing step): those are counted and reported separately, never silently.
The beam test does the same for `cor_correct_beam_host` with weights of variant "P" (peaked softmax, like a
trained model's) at the class-default beam parameters, so that tens of hypotheses per line are live, finals
are found among them and the search exits early -- while the hypothesis rows of all lines share the pair kernel.
"""
import numpy as np
import pytest
import torch

from oracle import decoders as od
from oracle.model import encode
from tests.common import make_model, synthetic_lines, topo_dict

pytestmark = pytest.mark.gpu

TOL_P = 1e-4
NEAR_TIE = 1e-4          # relative top-2 margin below which an argmax flip is a tie, not a parity failure


def _engine(topo, w, gemm_mode=0):
    from cor_asv_ann_b200.engine import Engine
    return Engine(topo_dict(topo), w, device=0, eos_index=1, gemm_mode=gemm_mode)


def c3v_batches(mapping, n_batches, seed=1013, batch=256):
    """lines of length 20..80 (+ newline), in reference batches of 256."""
    lines = synthetic_lines(mapping, n_batches * batch, 80, seed=seed, min_length=20)
    return [lines[b * batch:(b + 1) * batch] for b in range(n_batches)]


def _merge(mapping, batches):
    from cor_asv_ann_b200.vectorize import lines_to_sparse
    from cor_asv_ann_b200.pipeline import merge_batches
    encoded = [lines_to_sparse(mapping[0], lines, None) for lines in batches]
    return merge_batches(encoded)       # idx [B,Tmax,k], val, row_T (padded T of each row's batch), offsets


def _first_diff(a, b):
    n = min(len(a), len(b))
    for i in range(n):
        if a[i] != b[i]:
            return i
    return n


def test_greedy_c3v_one_call_pair_kernel():
    topo, w, mapping = make_model("c2")
    batches = c3v_batches(mapping, 17)
    idx, val, row_T, offsets = _merge(mapping, batches)
    B, T, _ = idx.shape
    assert B == 4352 and T == 81
    eng = _engine(topo, w, 0)
    eng.kernel_counters(reset=True)
    chars, probs, scores, lens, align = (None if t is None else t.numpy() for t in
                                         eng.correct_greedy_host(idx, val, row_T, want_align=True))
    kinds = eng.kernel_counters(reset=True)
    # the timed kernel: every LSTM GEMM of this call ran on CTA pairs, none on another variant
    assert kinds["tc_lstm_pair"] > 0 and kinds["tc_lstm_single"] == 0 and kinds["tc_lstm_narrow"] == 0 and kinds["ffma_lstm"] == 0, kinds
    assert kinds["tc_store_pair"] > 0, kinds
    text = [''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) for j in range(B)]

    # ---- (1) two whole reference batches against the oracle
    ties = real = 0
    worst_p = worst_a = 0.0
    for b in (0, 16):
        lines = batches[b]
        x = od.vectorize_lines(mapping, topo.voc_size, lines)
        trace = []
        r_text, r_probs, r_score, r_align = od.decode_batch_greedy(topo, w, mapping, x, trace=trace)
        Tb = x.shape[1]
        assert (row_T[offsets[b]:offsets[b + 1]] == Tb).all()
        for j in range(len(lines)):
            g = offsets[b] + j
            if text[g] != r_text[j]:
                s = _first_diff(text[g], r_text[j])
                row = np.sort(trace[s][j, 1:])[::-1]
                if (row[0] - row[1]) <= NEAR_TIE * row[0]:
                    ties += 1
                else:
                    real += 1
                    print("line %d of batch %d differs at step %d: margin %.3g" % (j, b, s, (row[0] - row[1]) / row[0]))
                n = s
            else:
                n = int(lens[g])
                assert abs(scores[g] - r_score[j]) < 1e-4
            if n:
                worst_p = max(worst_p, float(np.abs(probs[g, :n] - np.asarray(r_probs[j][:n], np.float32)).max()))
                ref_al = np.stack(r_align[j][:n])
                worst_a = max(worst_a, float(np.abs(align[:n, g, :Tb] - ref_al).max()))
    print("C3v vs oracle (512 lines, pair kernel): max |dp| %.3g, max |da| %.3g, %d near-tie string differences, %d real"
          % (worst_p, worst_a, ties, real))
    assert real == 0
    assert ties <= 5
    assert worst_p < TOL_P and worst_a < TOL_P

    # ---- (2) every row against the <= 512-row path (single-CTA kernel; pinned by the other tests)
    same = 0
    worst = 0.0
    for b0 in range(0, 17, 2):
        sub = batches[b0:b0 + 2]
        i2, v2, t2, _ = _merge(mapping, sub)
        c2, p2, s2, l2 = (t.numpy() for t in eng.correct_greedy_host(i2, v2, t2)[:4])
        for j in range(i2.shape[0]):
            g = offsets[b0] + j
            n = int(l2[j])
            if n == lens[g] and (c2[j, :n] == chars[g, :n]).all():
                same += 1
                if n:
                    worst = max(worst, float(np.abs(p2[j, :n] - probs[g, :n]).max()))
    kinds = eng.kernel_counters()
    assert kinds["tc_lstm_single"] > 0 and kinds["tc_lstm_pair"] == 0, kinds
    print("pair kernel vs single-CTA kernel: %d of %d strings identical, max |dp| %.3g" % (same, B, worst))
    assert same >= B - 8, same
    assert worst < TOL_P


def test_beam_competitive_one_call_pair_kernel():
    """cor_correct_beam_host over 512 lines of 20..60 characters at the C2 topology with peaked weights (variant
    "P") and the class-default beam parameters (seq2seq.py:160-169): the input character competes with the
    network's candidates, tens of live hypotheses per line and iteration, all lines' hypothesis rows in the same
    pair-kernel launches.  Eight lines are checked against the oracle (strings, statistics incl. the pop-order
    checksum, probabilities), every line against 64-line calls."""
    from cor_asv_ann_b200.vectorize import lines_to_sparse
    from tests.make_golden import stat_row
    topo, w, mapping = make_model("c2", variant="P")
    lines = synthetic_lines(mapping, 512, 60, seed=2013, min_length=20)
    params = dict(batch_size=256, beam_width_in=15, beam_threshold_in=0.2, beam_width_out=16, rejection_threshold=0.3)
    idx, val = lines_to_sparse(mapping[0], lines, None)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    eng = _engine(topo, w, 0)
    bp = eng.beam_params(**params)
    eng.kernel_counters(reset=True)
    chars, probs, scores, lens, _, stats = (None if t is None else t.numpy().copy() for t in
                                            eng.correct_beam_host(idx, val, row_T, bp, budget_bytes=24 << 30))
    kinds = eng.kernel_counters(reset=True)
    assert kinds["tc_lstm_pair"] > 0 and kinds["ffma_lstm"] == 0, kinds
    assert (stats[:, 0] >= 0).all()
    hyp_per_iter = stats[:, 1].sum() / max(stats[:, 0].sum(), 1)
    print("competitive beam: %.1f hypotheses per line and iteration, %.0f hypothesis-steps per line" % (hyp_per_iter, stats[:, 1].mean()))
    assert hyp_per_iter > 3.0          # a real search, not the one-hypothesis bracket
    text = [''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) for j in range(B)]

    # ---- oracle on a few lines (the batch encoding is part of the semantics: same padded T)
    x = od.vectorize_lines(mapping, topo.voc_size, lines)
    pick = [0, 1, 2, 3, 255, 256, 510, 511]
    enc = encode(topo, w, x[pick])
    bad = tie_explained = 0
    for q, j in enumerate(pick):
        st = {}
        try:
            r = next(od.decode_sequence_beam(topo, w, mapping, x[j], [e[q:q + 1] for e in enc], stats=st, **params))
        except StopIteration:
            r = None
        want = stat_row(st)
        same = (list(stats[j, :5]) == want[:5] and stats[j, 6] == want[6] and stats[j, 7] == want[7]
                and (r is None) == (lens[j] == 0) and (r is None or text[j] == r[0]))
        if not same:
            # a search that differs from the oracle's is a parity failure UNLESS either side logged a near-tie
            # (ordering decision within 1e-6 relative), which both arithmetics may resolve differently
            if stats[j, 5] > 0 or want[5] > 0:
                tie_explained += 1
            else:
                bad += 1
            print("line %d: stats %s vs oracle %s" % (j, list(stats[j]), want))
            continue
        if r is not None:
            n = int(lens[j])
            assert abs(scores[j] - r[2]) < 1e-4
            assert np.abs(probs[j, :n] - np.asarray(r[1], np.float32)).max() < TOL_P
    print("competitive beam vs oracle: %d of %d lines identical (strings, statistics, pop-order checksum), %d differ with a logged near-tie, %d without"
          % (len(pick) - bad - tie_explained, len(pick), tie_explained, bad))
    assert bad == 0
    assert tie_explained <= 2

    # ---- every line against 64-line calls: there the encoder is the persistent recurrence (mma.sync, another
    # summation order), so the two runs are two fp32-grade arithmetics, like the product and the oracle: a search
    # may differ only where one of them logged a near-tie
    diff_text = diff_search = unexplained = 0
    for b0 in range(0, B, 64):
        sl = slice(b0, b0 + 64)
        c2, p2, s2, l2, _, st2 = (None if t is None else t.numpy() for t in
                                  eng.correct_beam_host(idx[sl], None, row_T[sl], bp, budget_bytes=8 << 30))
        for j in range(64):
            g = b0 + j
            n = int(l2[j])
            same_text = n == lens[g] and (c2[j, :n] == chars[g, :n]).all()
            same_search = all(st2[j, c] == stats[g, c] for c in (0, 1, 2, 3, 4, 6, 7))
            diff_text += not same_text
            diff_search += not same_search
            if not (same_text and same_search) and st2[j, 5] == 0 and stats[g, 5] == 0:
                unexplained += 1
    print("512-line call vs 64-line calls: %d of %d strings differ, %d searches differ, %d of them without a logged near-tie"
          % (diff_text, B, diff_search, unexplained))
    assert unexplained <= B // 100         # decisions the near-tie log does not cover (e.g. the fringe truncation boundary)
    assert diff_text <= B // 20

"""GPU parity of the forward pass against the oracle: encoder, one decoder step, batch greedy.

All calls go through the C ABI (cor_asv_ann_b200.engine.Engine -> libcor_b200.so).
Tolerances: probabilities and alignments 1e-4 abs (BASELINE.json north_star, fp32 path);
hidden states 2e-5 abs per element (they are inputs to the former).
"""
import numpy as np
import pytest
import torch

from oracle import model as om
from oracle import decoders as od
from tests.common import make_model, synthetic_lines, topo_dict

pytestmark = pytest.mark.gpu

TOL_P = 1e-4
TOL_H = 5e-5


def _engine(topo, w, gemm_mode=0):
    from cor_asv_ann_b200.engine import Engine
    return Engine(topo_dict(topo), w, device=0, eos_index=1, gemm_mode=gemm_mode)


def _sparse(mapping, lines, conf=None):
    from cor_asv_ann_b200.vectorize import lines_to_sparse
    return lines_to_sparse(mapping[0], lines, conf)


def _dev(x, dtype):
    return None if x is None else torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def _encode_both(name, lines, gemm_mode=0, conf=None):
    topo, w, mapping = make_model(name)
    x = od.vectorize_lines(mapping, topo.voc_size, lines, conf)
    ref = om.encode(topo, w, x)
    eng = _engine(topo, w, gemm_mode)
    idx, val = _sparse(mapping, lines, conf)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    enc_out, states = eng.encode(_dev(idx, torch.int32), _dev(val, torch.float32), _dev(row_T, torch.int32))
    return topo, w, mapping, x, ref, eng, enc_out, states


@pytest.mark.parametrize("name", ["c1", "c2s", "c3res", "c4s"])
@pytest.mark.parametrize("gemm_mode", [1, 2])
def test_encoder_parity(name, gemm_mode):
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, 9, 23, seed=3, min_length=5) + ['']
    topo, w, mapping, x, ref, eng, enc_out, states = _encode_both(name, lines, gemm_mode)
    got = enc_out.permute(1, 0, 2).cpu().numpy()
    assert got.shape == ref[0].shape
    assert np.abs(got - ref[0]).max() < TOL_H
    st = states.cpu().numpy()
    for k in range(2 * topo.depth):
        assert np.abs(st[k] - ref[1 + k]).max() < TOL_H, "state %d" % k


def test_encoder_confidence_and_confmat_input():
    topo, w, mapping = make_model("c1")
    lines = synthetic_lines(mapping, 4, 12, seed=11)
    rng = np.random.default_rng(0)
    conf = [[float(rng.uniform(0.3, 1.0)) for _ in s] for s in lines]
    _, _, _, _, ref, _, enc_out, _ = _encode_both("c1", lines, 1, conf)
    assert np.abs(enc_out.permute(1, 0, 2).cpu().numpy() - ref[0]).max() < TOL_H
    # confusion network: chunks with alternatives of different length
    i2c = mapping[1]
    cm = [[[(i2c[5] + i2c[6], 0.6), (i2c[7], 0.3)], [], [(i2c[8], 0.9), (i2c[9], 0.05), (i2c[8], 0.8)], [('\n', 1.0)]],
          [[(i2c[20], 0.5)], [('\n', 0.7)]]]
    _, _, _, _, ref, _, enc_out, _ = _encode_both("c1", cm, 1, cm)
    assert np.abs(enc_out.permute(1, 0, 2).cpu().numpy() - ref[0]).max() < TOL_H


@pytest.mark.parametrize("gemm_mode", [1, 2])
@pytest.mark.parametrize("name", ["c1", "c2s", "c4s"])
def test_decoder_step_parity(name, gemm_mode):
    """cor_decode_step on the exact FFMA path (1) and on the tcgen05 three-pass path (2), incl. the LM second
    evaluation of the top cell (c4s)."""
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, 6, 19, seed=5, min_length=8)
    topo, w, mapping, x, ref, eng, enc_out, states = _encode_both(name, lines, gemm_mode)
    eng.kernel_counters(reset=True)
    B, T = x.shape[0], x.shape[1]
    d = topo.depth
    st = list(ref[1:])
    p = np.zeros((B, topo.voc_size), np.float32)
    for step in range(4):
        p_ref, lm_ref, st_ref = om.decode_step(topo, w, p, ref[0], st)
        h_in = torch.as_tensor(np.stack([st[2 * n] for n in range(d)]), device="cuda")
        c_in = torch.as_tensor(np.stack([st[2 * n + 1] for n in range(d)]), device="cuda")
        p_out, lm_out, h_out, c_out, a_out = eng.decode_step(
            torch.as_tensor(p, device="cuda"), h_in, c_in, torch.as_tensor(st[-1], device="cuda"))
        assert np.abs(p_out.cpu().numpy() - p_ref).max() < TOL_P
        assert np.abs(a_out.cpu().numpy() - st_ref[-1]).max() < TOL_P
        for n in range(d):
            assert np.abs(h_out[n].cpu().numpy() - st_ref[2 * n]).max() < TOL_H
            assert np.abs(c_out[n].cpu().numpy() - st_ref[2 * n + 1]).max() < TOL_H
        if topo.lm_predict:
            assert np.abs(lm_out.cpu().numpy() - lm_ref).max() < TOL_P
        p, st = p_ref, st_ref
    kinds = eng.kernel_counters()
    if gemm_mode == 2:
        assert kinds["ffma_lstm"] == 0 and kinds["tc_lstm_narrow"] + kinds["tc_lstm_single"] > 0, kinds
    else:
        assert kinds["ffma_lstm"] > 0 and kinds["tc_lstm_narrow"] + kinds["tc_lstm_single"] + kinds["tc_lstm_pair"] == 0, kinds


def test_decoder_step_broadcast_line():
    """Hypotheses of one line share its encoder output (seq2seq.py:1428-1429)."""
    topo, w, mapping = make_model("c1")
    lines = synthetic_lines(mapping, 3, 15, seed=8)
    topo, w, mapping, x, ref, eng, enc_out, states = _encode_both("c1", lines, 1)
    rng = np.random.default_rng(1)
    R, line = 3, 2
    p = rng.dirichlet(np.ones(topo.voc_size), size=R).astype(np.float32)
    st = [np.repeat(s[line:line + 1], R, axis=0) + rng.normal(0, .01, (R,) + s.shape[1:]).astype(np.float32) for s in ref[1:-1]]
    a = np.zeros((R, x.shape[1]), np.float32)
    a[:, 3] = 0.5
    a[:, 4] = 0.5
    p_ref, _, st_ref = om.decode_step(topo, w, p, ref[0][line:line + 1], st + [a])
    row_line = torch.full((R,), line, dtype=torch.int32, device="cuda")
    p_out, _, h_out, c_out, a_out = eng.decode_step(
        torch.as_tensor(p, device="cuda"), torch.as_tensor(np.stack(st[0::2]), device="cuda"),
        torch.as_tensor(np.stack(st[1::2]), device="cuda"), torch.as_tensor(a, device="cuda"), row_line)
    assert np.abs(p_out.cpu().numpy() - p_ref).max() < TOL_P
    assert np.abs(a_out.cpu().numpy() - st_ref[-1]).max() < TOL_P


def _greedy_compare(name, lines, gemm_mode, conf=None):
    topo, w, mapping = make_model(name)
    trace = []
    x = od.vectorize_lines(mapping, topo.voc_size, lines, conf)
    r_text, r_probs, r_score, r_align = od.decode_batch_greedy(topo, w, mapping, x, trace=trace)
    eng = _engine(topo, w, gemm_mode)
    idx, val = _sparse(mapping, lines, conf)
    B, T, _ = idx.shape
    chars, probs, scores, lens, align = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32), want_align=True)
    chars, probs, scores, lens, align = (t.numpy() for t in (chars, probs, scores, lens, align))
    mismatched = 0
    for j in range(B):
        text = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]])
        if text != r_text[j]:
            mismatched += 1
            continue
        n = lens[j]
        assert n == len(r_probs[j])
        if n:
            assert np.abs(probs[j, :n] - np.array(r_probs[j])).max() < TOL_P
            assert abs(scores[j] - r_score[j]) < 1e-4
            ref_al = np.stack(r_align[j])
            assert np.abs(align[:n, j, :] - ref_al).max() < TOL_P
    return mismatched, B


@pytest.mark.parametrize("name", ["c1", "c2s", "c3res"])
@pytest.mark.parametrize("gemm_mode", [1, 2])
def test_batch_greedy_parity(name, gemm_mode):
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, 11, 21, seed=13, min_length=6) + ['']
    bad, B = _greedy_compare(name, lines, gemm_mode)
    assert bad == 0, "%d of %d strings differ" % (bad, B)


def test_batch_greedy_c2_size():
    """Full C2/C3 topology (depth 2, width 512, V 512, T 81) on a few lines."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, 12, 80, seed=1003)
    for mode in (1, 2):
        bad, B = _greedy_compare("c2", lines, mode)
        assert bad == 0, "gemm_mode %d: %d of %d strings differ" % (mode, bad, B)


@pytest.mark.parametrize("gemm_mode", [1, 2])
def test_two_reference_batches_in_one_call(gemm_mode):
    """Rows of reference batches with different padded T share one call (row_T); gemm_mode 2 with <= 32 rows
    runs the encoder as the persistent recurrence (recur.cu)."""
    topo, w, mapping = make_model("c2s")
    a = synthetic_lines(mapping, 5, 14, seed=21, min_length=4)
    b = synthetic_lines(mapping, 4, 25, seed=22, min_length=9)
    ra = od.correct_lines(topo, w, mapping, a, fast=True)
    rb = od.correct_lines(topo, w, mapping, b, fast=True)
    eng = _engine(topo, w, gemm_mode)
    idx, _ = _sparse(mapping, a + b)
    Ta, Tb = max(map(len, a)), max(map(len, b))
    row_T = np.array([Ta] * len(a) + [Tb] * len(b), np.int32)
    chars, probs, scores, lens, _ = (t.numpy() for t in eng.correct_greedy_host(idx, None, row_T)[:4] + (torch.zeros(1),))
    got = [''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) for j in range(len(a) + len(b))]
    assert got == ra[0] + rb[0]
    assert np.allclose(scores, np.array(ra[2] + rb[2]), atol=1e-4)


def test_small_magnitude_weights_variant_k():
    """Keras-init weights (variant K: embedding N(0, 0.001), zero-noise biases): every element of E, of the derived table
    PE = E.Kx and of the logits operand has a SUBNORMAL fp16 residual (epilogue.cuh:8-12), so the tensor path is exact
    only in absolute terms there.  End to end that is still far inside the parity budget: tensor path (gemm_mode 2)
    against the exact FFMA path (gemm_mode 1) and against the oracle -- probabilities within 1e-4 abs (they are ~1/V, so
    the check is also made relative: 1e-3), alignments within 1e-4; the near-uniform softmax makes near-tie argmaxes
    common, and those are logged."""
    topo, w, mapping = make_model("c2s", variant="K")
    lines = synthetic_lines(mapping, 24, 18, seed=77, min_length=6)
    x = od.vectorize_lines(mapping, topo.voc_size, lines, None)
    r_text, r_probs, _, r_align = od.decode_batch_greedy(topo, w, mapping, x)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    res = {}
    for mode in (1, 2):
        eng = _engine(topo, w, mode)
        out = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32), want_align=True)
        res[mode] = [t.numpy().copy() for t in out] + [eng.last_ties.numpy().copy()]
    for mode in (1, 2):
        chars, probs, scores, lens, align, ties = res[mode]
        same, dp, rel, da, untied, edge, steps = 0, 0.0, 0.0, 0.0, [], 0, 0
        for j in range(B):
            text = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]])
            # the fast decoder feeds the softmax back, not the argmax: every step is comparable whatever the strings say
            n = min(int(lens[j]), len(r_probs[j]))
            k = 0
            while k < min(len(text), len(r_text[j])) and text[k] == r_text[j][k]:
                k += 1
            same += int(text == r_text[j])
            if k:       # probabilities belong to the emitted characters: compare where those agree
                ref_p = np.array(r_probs[j][:k])
                dp = max(dp, float(np.abs(probs[j, :k] - ref_p).max()))
                rel = max(rel, float((np.abs(probs[j, :k] - ref_p) / ref_p).max()))
            if n:
                # A near-uniform attention puts its centre (almost) exactly on an integer, where the inclusive window edge
                # |centre - j| <= 5 hangs on the last bit: steps whose window SUPPORT differs from the oracle's
                # are counted, the others compared.
                ga, ra = align[:n, j, :], np.stack(r_align[j][:n])
                ok = ((ga > 0) == (ra > 0)).all(axis=1)
                edge += int((~ok).sum()); steps += n
                if ok.any():
                    da = max(da, float(np.abs(ga[ok] - ra[ok]).max()))
            if text != r_text[j] and ties[j] == 0:
                untied.append(j)
        print("variant K, gemm_mode %d: %d of %d strings identical to the oracle, %d near-ties logged; max |dp| %.3g (relative %.3g), "
              "max |da| %.3g on %d steps, %d steps with another window support" % (mode, same, B, int(ties.sum()), dp, rel, da, steps, edge))
        assert dp < TOL_P and rel < 1e-3 and da < TOL_P, "mode %d" % mode
        assert edge <= steps // 4, "mode %d" % mode
        assert not untied, "mode %d: lines %s differ from the oracle without a logged near-tie" % (mode, untied)


# Stated tolerance of the reduced-precision variant (gemm_mode 3, single-pass fp16 tensor products; the north
# star's "bf16 GEMM variants get a stated looser tolerance").  Measured on 2 048 lines by scripts/measure_mode3.py
# (profiles/r01_mode3_drift.json: 99.1 % identical strings, max |dp| 3.9e-7 where strings agree, score drift 3.8e-5 --
# with random-init weights the probabilities themselves are ~1/V, so this is ~2e-4 relative).
TOL_P_SINGLE_PASS = 1e-5


def test_single_pass_variant_stated_tolerance():
    """gemm_mode 3 against the fp32-grade product path (gemm_mode 2, itself checked against the oracle above)
    at the C2/C3 topology: per-character probabilities within TOL_P_SINGLE_PASS up to the first differing
    character, and the two paths are really different kernels (the drift is not zero)."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, 256, 80, seed=4242)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    res = {}
    for mode in (2, 3):
        eng = _engine(topo, w, mode)
        chars, probs, scores, lens, _ = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32), want_align=False)
        res[mode] = tuple(t.numpy().copy() for t in (chars, probs, lens))
    (c2, p2, l2), (c3, p3, l3) = res[2], res[3]
    worst, identical = 0.0, 0
    for j in range(B):
        n = int(min(l2[j], l3[j]))
        d = np.nonzero(c2[j, :n] != c3[j, :n])[0]
        k = int(d[0]) if len(d) else n
        identical += int(len(d) == 0 and l2[j] == l3[j])
        if k:
            worst = max(worst, float(np.abs(p2[j, :k] - p3[j, :k]).max()))
    print("single pass: %d of %d strings identical, max |dp| on common prefixes %.3g" % (identical, B, worst))
    assert 0.0 < worst < TOL_P_SINGLE_PASS
    assert identical >= (B * 19) // 20


@pytest.mark.parametrize("n_lines", [1, 16, 17, 32, 33, 48, 64])
def test_page_sized_batches_c2(n_lines):
    """Page-sized calls at the C2/C3 topology (<= 64 rows in blocks of 32: persistent encoder recurrence with the weights
    resident in shared memory, one launch per layer) against the oracle: strings, probabilities, alignments."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, n_lines, 40, seed=900 + n_lines, min_length=12)
    bad, B = _greedy_compare("c2", lines, 2)
    assert bad == 0, "%d of %d strings differ" % (bad, B)


def test_page_path_agrees_with_large_batch_path():
    """Two kernel families, one answer: 512 lines decoded in ONE call (tcgen05 GEMM per timestep) and the same lines
    in 16 page-sized calls (persistent recurrence + fused page decoder, mma.sync, different summation orders).
    Both are fp32-grade, so the strings agree except at near-ties and the probabilities within the parity budget."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, 512, 60, seed=777, min_length=60)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    eng = _engine(topo, w, 2)
    row_T = np.full(B, T, np.int32)
    big = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T)[:4]]
    same, worst = 0, 0.0
    for b0 in range(0, B, 32):
        sl = slice(b0, b0 + 32)
        chars, probs, scores, lens = [t.numpy() for t in eng.correct_greedy_host(idx[sl], None if val is None else val[sl], row_T[sl])[:4]]
        for j in range(32):
            n = int(lens[j])
            if n == big[3][b0 + j] and (chars[j, :n] == big[0][b0 + j, :n]).all():
                same += 1
                if n:
                    worst = max(worst, float(np.abs(probs[j, :n] - big[1][b0 + j, :n]).max()))
    print("page path vs large-batch path: %d of %d strings identical, max |dp| %.3g" % (same, B, worst))
    assert same >= B - 5, same
    assert worst < TOL_P

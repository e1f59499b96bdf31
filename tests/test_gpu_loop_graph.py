"""The decoder loops as device-side WHILE graphs (csrc/graphs.h) against the host-driven loops and the oracle.

The graph path and the host loop launch the same kernels on the same buffers, so their results must be IDENTICAL
bit for bit; the graph path additionally stops as soon as every line has finished (the reference keeps stepping
finished lines and ignores them, seq2seq.py:1255-1256), which must not change anything either.
"""
import os

import numpy as np
import pytest
import torch

from oracle import decoders as od
from tests.common import BEAM_CASES, make_model, synthetic_lines, topo_dict

pytestmark = pytest.mark.gpu


def _engine(topo, w, gemm_mode=0):
    from cor_asv_ann_b200.engine import Engine
    return Engine(topo_dict(topo), w, device=0, eos_index=1, gemm_mode=gemm_mode)


def _sparse(mapping, lines):
    from cor_asv_ann_b200.vectorize import lines_to_sparse
    return lines_to_sparse(mapping[0], lines)


class _host_loop:
    """COR_NO_GRAPH=1 for the calls inside (read by the library at call time)."""
    def __enter__(self):
        os.environ["COR_NO_GRAPH"] = "1"
    def __exit__(self, *a):
        os.environ.pop("COR_NO_GRAPH", None)


def _ending_model(name, boost=4.0, recurrent=12.0):
    """A model whose lines END, at different steps, which a random-init model's never do (its decoder settles on a
    fixed point): peaked weights (variant P), the decoder's recurrent kernels scaled up (no fixed point any more) and
    the newline row of the tied embedding boosted, so the newline wins the argmax at some step of most lines.
    Such dynamics amplify rounding differences, so this model is compared between the two loop drivers of the
    SAME kernels (bit-identical by construction), not with the oracle."""
    topo, w, mapping = make_model(name, variant="P")
    w = {k: [a.copy() for a in v] for k, v in w.items()}
    w["char_input_projection"][0][1] *= np.float32(boost)
    for n in range(1, topo.depth + 1):
        i = 5 if n == topo.depth else 1
        w["decoder_lstm_%d" % n][i] = (w["decoder_lstm_%d" % n][i] * np.float32(recurrent)).astype(np.float32)
    return topo, w, mapping


@pytest.mark.parametrize("name,n_lines,length", [("c2s", 400, 20), ("c2", 400, 30)])
def test_greedy_graph_early_exit_identical(name, n_lines, length):
    """Fast greedy decode of lines that all end before 2T: graph == host loop bit for bit, and the graph ran fewer
    timesteps than the 2T the host loop always launches."""
    topo, w, mapping = _ending_model(name)
    lines = synthetic_lines(mapping, n_lines, length, seed=5151, min_length=length // 2)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    eng = _engine(topo, w, 2)
    with _host_loop():
        first = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32))[3].numpy().copy()
    keep = np.nonzero(first < T)[0]         # lines that end within half of the 2T steps
    assert len(keep) > 64, "the test model must end enough lines early (more than a page of them)"
    idx = np.ascontiguousarray(idx[keep]); val = None if val is None else np.ascontiguousarray(val[keep])
    B = len(keep)
    row_T = np.full(B, T, np.int32)
    eng.launch_count(reset=True)
    g = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    n_graph = eng.launch_count(reset=True)
    with _host_loop():
        h = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    n_host = eng.launch_count(reset=True)
    chars, probs, scores, lens, align = g
    assert (lens == first[keep]).all() and (lens == h[3]).all()
    for j in range(B):
        n = int(lens[j])
        assert (chars[j, :n] == h[0][j, :n]).all()
        assert (probs[j, :n] == h[1][j, :n]).all()
        assert (align[:n, j] == h[4][:n, j]).all()
    assert (scores == h[2]).all()
    assert n_graph < 0.8 * n_host, (n_graph, n_host)
    print("greedy graph: %d launches against %d of the host loop, longest line %d of %d steps, %d lines" % (n_graph, n_host, lens.max(), 2 * T, B))


def test_greedy_graph_row_compaction_mixed():
    """Lines that end at scattered steps NEXT TO lines that never do: finished rows are compacted out of the GEMMs while the
    others run their 2T steps; every line's result is what the host loop (all rows, all steps) gives, bit for bit --
    incl. the alignment history, which is indexed by line while rows move."""
    topo, w, mapping = _ending_model("c2")
    lines = synthetic_lines(mapping, 300, 24, seed=777, min_length=8) + ['']
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    row_T[100:200] = T - 5           # a second reference batch with a shorter padded length in the same call
    idx[100:200, T - 5:, :] = -1
    eng = _engine(topo, w, 2)
    g = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    with _host_loop():
        h = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    chars, probs, scores, lens, align = g
    ended = int((lens < 2 * row_T).sum())
    assert 30 < ended < B - 30, "the test needs a mix of ending and non-ending lines (%d of %d end)" % (ended, B)
    assert np.array_equal(lens, h[3]) and np.array_equal(scores, h[2])
    for j in range(B):
        n = int(lens[j])
        assert np.array_equal(chars[j, :n], h[0][j, :n]) and np.array_equal(probs[j, :n], h[1][j, :n])
        assert np.array_equal(align[:n, j], h[4][:n, j])
    assert np.array_equal(eng.last_ties.numpy(), eng.last_ties.numpy())


def test_greedy_graph_full_length_identical():
    """Random-init weights (no line ends): all 2T steps run inside the graph; identical to the host loop."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, 80, 25, seed=99, min_length=10)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    eng = _engine(topo, w, 2)
    g = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    with _host_loop():
        h = [t.numpy().copy() for t in eng.correct_greedy_host(idx, val, row_T, want_align=True)]
    for a, b in zip(g, h):
        assert np.array_equal(a, b)
    assert (g[3] == 2 * T).all()


@pytest.mark.parametrize("case", ["c2s_mix", "c2s_rej0", "c4s_lm_mix", "c1_mix_wide"])
def test_beam_graph_identical_to_host_loop(case):
    name, n, length, seed, params = BEAM_CASES[case]
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, n, length, seed=seed, min_length=max(3, length // 2))
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    eng = _engine(topo, w, 2)
    bp = eng.beam_params(**params)
    eng.launch_count(reset=True)
    g = [None if t is None else t.numpy().copy() for t in eng.correct_beam_host(idx, val, row_T, bp, want_align=True, budget_bytes=2 << 30)]
    n_graph = eng.launch_count(reset=True)
    with _host_loop():
        h = [None if t is None else t.numpy().copy() for t in eng.correct_beam_host(idx, val, row_T, bp, want_align=True, budget_bytes=2 << 30)]
    n_host = eng.launch_count(reset=True)
    for a, b in zip(g, h):
        assert np.array_equal(a, b)
    # kernels run: the graph stops with the last line (plus its one condition kernel), the host loop looks every 8 iterations
    assert 0 < n_graph <= n_host + 1, (n_graph, n_host)


def test_beam_graph_page_of_c2_lines():
    """One page (32 lines) at the shipped topology, the OCR-D default mode: graph == host loop, and == oracle on a few lines."""
    topo, w, mapping = make_model("c2")
    lines = synthetic_lines(mapping, 32, 30, seed=4711, min_length=12)
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    eng = _engine(topo, w, 2)
    params = dict(batch_size=256, beam_width_in=4, beam_threshold_in=0.2, beam_width_out=16, rejection_threshold=0.5)
    bp = eng.beam_params(**params)
    g = [None if t is None else t.numpy().copy() for t in eng.correct_beam_host(idx, val, row_T, bp, budget_bytes=2 << 30)]
    with _host_loop():
        h = [None if t is None else t.numpy().copy() for t in eng.correct_beam_host(idx, val, row_T, bp, budget_bytes=2 << 30)]
    for a, b in zip(g, h):
        assert np.array_equal(a, b)
    chars, probs, scores, lens = g[:4]
    ref_text, ref_probs, ref_score, _ = od.correct_lines(topo, w, mapping, lines[:3], fast=False, greedy=False, beam_params=params)
    for j in range(3):
        got = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) if lens[j] > 0 else lines[j]
        assert got == ref_text[j]


@pytest.mark.parametrize("case", ["c2s_mix", "c2s_rej0", "c1_mix_wide"])
def test_line_kernel_path_identical(case):
    """The opt-in line-kernel path of a page (COR_LINE_KERNEL=1: gather with the scan, one block per line for expand + push +
    pop) gives the results of the default path bit for bit."""
    name, n, length, seed, params = BEAM_CASES[case]
    topo, w, mapping = make_model(name)
    lines = synthetic_lines(mapping, n, length, seed=seed, min_length=max(3, length // 2))
    idx, val = _sparse(mapping, lines)
    B, T, _ = idx.shape
    row_T = np.full(B, T, np.int32)
    eng = _engine(topo, w, 2)
    bp = eng.beam_params(**params)
    ref = [None if t is None else t.numpy().copy() for t in eng.correct_beam_host(idx, val, row_T, bp, want_align=True, budget_bytes=2 << 30)]
    os.environ["COR_LINE_KERNEL"] = "1"
    try:
        eng2 = _engine(topo, w, 2)          # own graph cache: the key does not know the switch
        got = [None if t is None else t.numpy().copy() for t in eng2.correct_beam_host(idx, val, row_T, bp, want_align=True, budget_bytes=2 << 30)]
    finally:
        os.environ.pop("COR_LINE_KERNEL", None)
    # (separate engines = separate workspaces: alignment rows past a line's length are whatever the buffer held)
    chars, probs, scores, lens, align, stats = got
    for a, b in zip((chars, probs, scores, lens, stats), (ref[0], ref[1], ref[2], ref[3], ref[5])):
        assert np.array_equal(a, b)
    for j in range(B):
        assert np.array_equal(align[j, :lens[j]], ref[4][j, :lens[j]])

"""Row N1: the device Viterbi re-alignment (cor_alignment_paths) against the oracle restatement of
wrapper/transcode.py:279-349 and against outputs of the reference's own function
(tests/golden/ref_alignment2path.npz).  Bit-exact: integer paths and the fp64 distance."""
import os

import numpy as np
import pytest

from oracle import alignment_path as oap
from oracle import decoders as od
from tests import ref_harness as rh
from tests.common import make_model, synthetic_lines, topo_dict

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from cor_asv_ann_b200.engine import Engine
    topo, w, _ = make_model("c1")
    return Engine(topo_dict(topo), w, device=0)


def _random_case(rng, S, T, n):
    a = rng.random((S, T)).astype(np.float32) ** 4
    a /= a.sum(1, keepdims=True)
    if n % 3 == 0:
        a = (np.round(a * 4) / 4).astype(np.float32)            # exact ties
    if n % 5 == 0:
        a[rng.integers(0, S)] = 0                               # an all-zero step
    return a.astype(np.float32)


@pytest.mark.parametrize("layout", ["greedy", "beam"])
def test_paths_match_oracle_random(eng, layout):
    import torch
    rng = np.random.default_rng(23)
    B, S, T = 96, 40, 27
    a = np.stack([_random_case(rng, S, T, n) for n in range(B)])              # [B,S,T]
    i_max = rng.integers(1, T + 1, size=B).astype(np.int32)
    j_max = rng.integers(1, S + 1, size=B).astype(np.int32)
    i_max[:4], j_max[:4] = [1, T, 1, T], [1, 1, S, S]
    dev_a = torch.as_tensor(a if layout == "beam" else a.transpose(1, 0, 2).copy(), device="cuda")
    path, dist = eng.alignment_paths(dev_a, layout, i_max, j_max, min_score=1.0 / 80)
    path, dist = path.cpu().numpy(), dist.cpu().numpy()
    for b in range(B):
        want_path, want_dist = oap.alignment_path(a[b], int(i_max[b]), int(j_max[b]), 1.0 / 80)
        assert (path[b, :i_max[b] + 1] == want_path).all(), b
        assert (path[b, i_max[b] + 1:] == -1).all()
        assert dist[b] == want_dist, b


def test_paths_match_reference_fixture(eng):
    import torch
    g = np.load(os.path.join(rh.GOLDEN, "ref_alignment2path.npz"), allow_pickle=False)
    for k in range(int(g["n"][0])):
        a = g["a%d" % k]
        i_max, j_max = (int(v) for v in g["dims%d" % k])
        path, dist = eng.alignment_paths(torch.as_tensor(a[None], device="cuda"), "beam", [i_max], [j_max],
                                         min_score=float(g["ms%d" % k][0]))
        assert (path.cpu().numpy()[0, :i_max + 1] == g["path%d" % k]).all(), k
        # the fixture's distance was accumulated in float32 (numpy 2 in the build container); the
        # product accumulates in float64 like the reference's pinned numpy
        assert abs(float(dist[0]) - float(g["dist%d" % k][0])) < 1e-4


def test_identity_alignment_and_degenerate_rows(eng):
    import torch
    a = torch.zeros((3, 6, 5), device="cuda")
    path, dist = eng.alignment_paths(a, "beam", [5, 0, 4], [-5, 3, 0])
    path, dist = path.cpu().numpy(), dist.cpu().numpy()
    want, wdist = oap.alignment_path(np.eye(5, dtype=np.float32), 5, 5, 1.0 / 70)
    assert (path[0] == want).all() and dist[0] == wdist == 0.0
    assert path[1, 0] == 0 and dist[1] == 0 and path[2, 0] == 0 and path[2, 4] == 0


@pytest.mark.parametrize("mode", ["fast", "greedy", "beam"])
def test_correct_lines_paths(mode, tmp_path):
    """The host class: same strings / probabilities / scores as correct_lines, and the path the
    reference's wrapper would compute from the returned soft alignments."""
    from tests.test_gpu_seq2seq import _loaded
    topo, w, mapping = make_model("c2s")
    lines = synthetic_lines(mapping, 7, 16, seed=97, min_length=6)
    s = _loaded(topo, w, mapping, tmp_path)
    s.rejection_threshold, s.beam_width_in, s.batch_size = 0.1, 4, 32
    fast, greedy = mode == "fast", mode != "beam"
    ref = s.correct_lines(lines, fast=fast, greedy=greedy)
    got = s.correct_lines_paths(lines, fast=fast, greedy=greedy)
    assert got[0] == ref[0]
    assert [[float(p) for p in pl] for pl in got[1]] == [[float(p) for p in pl] for pl in ref[1]]
    assert [float(x) for x in got[2]] == [float(x) for x in ref[2]]
    for j, line in enumerate(lines):
        want_path, want_dist = oap.alignment_path(ref[3][j], len(line), len(ref[0][j]), 1.0 / topo.voc_size)
        d = got[3][j]
        assert d == {int(i): int(v) for i, v in enumerate(want_path) if v >= 0}
        assert got[4][j] == want_dist

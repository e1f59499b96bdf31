"""The PRODUCT (CUDA path behind the drop-in Sequence2Sequence class) against outputs of the
reference's own decoder code (tests/golden/ref_*.npz, made by tests/ref_harness.py)."""
import os

import numpy as np
import pytest

from tests import ref_harness as rh
from tests.test_gpu_seq2seq import _loaded

pytestmark = pytest.mark.gpu
ALL_CASES = list(rh.DECODER_CASES) + list(rh.KANT_CASES)


@pytest.mark.parametrize("name", ALL_CASES)
def test_product_matches_reference_fixture(name, tmp_path):
    g = np.load(os.path.join(rh.GOLDEN, "ref_%s.npz" % name), allow_pickle=False)
    topo, w, mapping, lines, mode, params = rh.case_inputs(name)
    s = _loaded(topo, w, mapping, tmp_path)
    for k, v in params.items():
        setattr(s, k, v)
    conf = lines if name in rh.KANT_CASES else None
    res = s.correct_lines(lines, conf, fast=(mode == "fast"), greedy=(mode != "beam"))
    got = rh.pack_result(res)
    assert [t.decode("utf-8") for t in g["texts"]] == [t.decode("utf-8") for t in got["texts"]]
    assert (g["lens"] == got["lens"]).all()
    assert np.abs(g["probs"] - got["probs"]).max(initial=0) < 1e-4          # north-star tolerance, fp32
    assert np.abs(g["scores"] - got["scores"]).max(initial=0) < 1e-4
    assert np.abs(g["align_centre"] - got["align_centre"]).max(initial=0) < 2e-3

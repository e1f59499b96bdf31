"""How far the single-pass fp16 variant (gemm_mode 3) drifts from the fp32-grade product path (gemm_mode 2):
string agreement and probability differences of the fast batch-greedy decoder at the C2/C3 topology.
Writes gpurun_out/mode3.json; the numbers set the tolerance stated in DESIGN.md section 3 and asserted by
tests/test_gpu_core.py::test_single_pass_variant_stated_tolerance."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.common import make_model, synthetic_lines, topo_dict  # noqa: E402
from cor_asv_ann_b200.engine import Engine  # noqa: E402
from cor_asv_ann_b200.vectorize import lines_to_sparse  # noqa: E402


def run(variant, n, length):
    topo, w, mapping = make_model("c2", variant=variant)
    lines = synthetic_lines(mapping, n, length, seed=4242)
    idx, val = lines_to_sparse(mapping[0], lines)
    B, T, _ = idx.shape
    res = {}
    for mode in (2, 3):
        eng = Engine(topo_dict(topo), w, device=0, eos_index=1, gemm_mode=mode)
        chars, probs, scores, lens, _ = eng.correct_greedy_host(idx, val, np.full(B, T, np.int32), want_align=False)
        res[mode] = tuple(t.numpy().copy() for t in (chars, probs, scores, lens))
        del eng
    c2, p2, s2, l2 = res[2]
    c3, p3, s3, l3 = res[3]
    same = [j for j in range(B) if l2[j] == l3[j] and (c2[j, :l2[j]] == c3[j, :l3[j]]).all()]
    first_diff = []
    for j in range(B):
        n_ = min(l2[j], l3[j])
        d = np.nonzero(c2[j, :n_] != c3[j, :n_])[0]
        first_diff.append(int(d[0]) if len(d) else int(n_))
    dp = max((np.abs(p2[j, :l2[j]] - p3[j, :l2[j]]).max() for j in same if l2[j]), default=0.0)
    # probabilities up to the first differing character on every line (before the feedback diverges)
    dp_prefix = max((np.abs(p2[j, :first_diff[j]] - p3[j, :first_diff[j]]).max() for j in range(B) if first_diff[j]), default=0.0)
    return {"variant": variant, "lines": B, "length": length, "identical_strings": len(same),
            "max_abs_dprob_identical_lines": float(dp), "max_abs_dprob_common_prefix": float(dp_prefix),
            "max_abs_dscore_identical_lines": float(max((abs(s2[j] - s3[j]) for j in same), default=0.0)),
            "mean_common_prefix": float(np.mean(first_diff)), "mean_len": float(np.mean(l2))}


if __name__ == "__main__":
    out = [run("U", 2048, 80), run("K", 512, 80)]
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/mode3.json", "w"), indent=1)
    print(json.dumps(out))

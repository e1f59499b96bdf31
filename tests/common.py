"""Shared helpers for the tests: synthetic alphabets, lines, topologies."""
import numpy as np

from oracle.model import Topology, init_weights


def synthetic_alphabet(V):
    """index 0 '', index 1 '\\n', then code points ascending from U+0020 skipping U+007F-U+00A0."""
    chars = ['', '\n']
    cp = 0x20
    while len(chars) < V:
        if not (0x7F <= cp <= 0xA0):
            chars.append(chr(cp))
        cp += 1
    return ({c: i for i, c in enumerate(chars)}, {i: c for i, c in enumerate(chars)})


def synthetic_lines(mapping, n, length, seed, min_length=None):
    """n lines of `length` chars i.i.d. uniform over indices 2..V-1, plus '\\n' (lengths uniform in
    [min_length, length] when given)."""
    rng = np.random.default_rng(seed)
    V = len(mapping[0])
    out = []
    for _ in range(n):
        L = length if min_length is None else int(rng.integers(min_length, length + 1))
        out.append(''.join(mapping[1][int(i)] for i in rng.integers(2, V, size=L)) + '\n')
    return out


CONFIGS = {
    # name: (Topology kwargs, V)
    "c1": dict(width=128, depth=1, voc_size=70),
    "c2s": dict(width=128, depth=2, voc_size=96),          # small stand-in for depth 2
    "c2": dict(width=512, depth=2, voc_size=512),
    "c3res": dict(width=64, depth=3, voc_size=80, residual_connections=True, bridge_dense=True),
    "c4s": dict(width=64, depth=3, voc_size=80, deep_bidirectional_encoder=True,
                residual_connections=True, lm_predict=True),
}


def make_model(name, seed=7, variant="U"):
    topo = Topology(**CONFIGS[name])
    return topo, init_weights(topo, seed, variant), synthetic_alphabet(topo.voc_size)


def topo_dict(topo):
    return dict(width=topo.width, depth=topo.depth, voc_size=topo.voc_size,
                residual_connections=topo.residual_connections,
                deep_bidirectional_encoder=topo.deep_bidirectional_encoder,
                bridge_dense=topo.bridge_dense, lm_predict=topo.lm_predict,
                window_width=topo.window_width)


# ---------------------------------------------------------------------------- fixtures
import json as _json
import os as _os

GOLDEN_DIR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden")

# beam known-answer cases: name -> (topology, lines, max length, seed, decode_sequence_beam parameters)
BEAM_CASES = {
    "c2s_rej05": ("c2s", 6, 14, 31, dict(rejection_threshold=0.5)),
    "c2s_rej0": ("c2s", 3, 9, 32, dict(rejection_threshold=0, batch_size=32, beam_width_in=5)),
    "c2s_rej0_wide": ("c2s", 2, 7, 36, dict(rejection_threshold=0, batch_size=256, beam_width_in=15)),
    "c1_rej01": ("c1", 5, 12, 33, dict(rejection_threshold=0.1, batch_size=64)),
    "c3res_rej03": ("c3res", 4, 11, 34, dict(rejection_threshold=0.3, batch_size=48, beam_width_in=6)),
    "c4s_lm": ("c4s", 4, 10, 35, dict(rejection_threshold=0.3, batch_size=32, beam_width_in=4)),
    # low rejection thresholds: the input character competes with the network's candidates, several
    # hypotheses per iteration, finals found among them, early exit by beam_width_out
    "c2s_mix": ("c2s", 3, 9, 32, dict(rejection_threshold=0.03, batch_size=64, beam_width_in=5)),
    "c1_mix": ("c1", 3, 9, 32, dict(rejection_threshold=0.05, batch_size=64, beam_width_in=5)),
    "c1_mix_wide": ("c1", 3, 9, 32, dict(rejection_threshold=0.05, batch_size=256, beam_width_in=15)),
    "c1_mix_narrow": ("c1", 3, 9, 32, dict(rejection_threshold=0.03, batch_size=16, beam_width_in=3)),
    "c4s_lm_mix": ("c4s", 3, 9, 37, dict(rejection_threshold=0.05, batch_size=32, beam_width_in=4)),
}


def kant_fixture():
    """(alphabet, pages) of tests/golden/kant_lines.json (made by tests/make_golden.py)."""
    with open(_os.path.join(GOLDEN_DIR, "kant_lines.json"), encoding="utf-8") as f:
        d = _json.load(f)
    return d["alphabet"], d["pages"]


def kant_model(alphabet, seed=11):
    """Config 1: depth 1 / width 128 over the alphabet of the kant lines."""
    topo = Topology(width=128, depth=1, voc_size=len(alphabet))
    mapping = ({c: i for i, c in enumerate(alphabet)}, {i: c for i, c in enumerate(alphabet)})
    return topo, init_weights(topo, seed, "U"), mapping

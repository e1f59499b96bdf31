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

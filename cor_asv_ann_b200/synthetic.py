"""Synthetic alphabets, lines and random-init weights of a named topology: what
``bench.py``, ``smoke()`` and the scripts feed the engine when there is no trained model and no dataset.
Product side: nothing here touches ``oracle/``; the oracle takes the same weight dictionaries as input.

Weight variants:
  "K"  Keras default initialisers (ocrd_cor_asv_ann/lib/seq2seq.py:240: embedding N(0, 0.001); glorot-uniform
       kernels; orthogonal recurrent kernels; zero biases with unit forget gate) -- near-uniform softmax;
  "U"  as K with embedding rows of unit norm (what the regulariser at seq2seq.py:549-551 drives trained models
       to) and N(0, 0.1) noise on the LSTM biases, so that padding steps change the state;
  "P"  as U with embedding rows of norm 12: logits of standard deviation ~3 like a trained model's, a selective
       softmax -- the variant on which the beam search is a search (tens of live hypotheses per line).
  "E"  as P with the decoder's recurrent kernels scaled by 12 (the decoder state no longer settles on a fixed point)
       and the newline row of the embedding scaled by 4: lines END, at different steps, which no random-init model's
       ever do -- the variant on which the fast decoder's early exit shows (it amplifies rounding differences, so
       it is compared between drivers of the same kernels, not with the oracle).
"""
import numpy as np


def synthetic_alphabet(voc_size):
    """(char -> index, index -> char): index 0 '', index 1 newline, then code points ascending from U+0020
    skipping U+007F-U+00A0."""
    chars = ['', '\n']
    cp = 0x20
    while len(chars) < voc_size:
        if not 0x7F <= cp <= 0xA0:
            chars.append(chr(cp))
        cp += 1
    return {c: i for i, c in enumerate(chars)}, {i: c for i, c in enumerate(chars)}


def line_indices(n_lines, length, voc_size, seed, min_length=None):
    """int32 [n, length + 1, 1]: `length` characters uniform over indices 2..V-1, then the newline (index 1);
    with `min_length`, lengths uniform in [min_length, length] and -1 (true zero) after the newline."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(2, voc_size, size=(n_lines, length + 1, 1), dtype=np.int32)
    if min_length is None:
        idx[:, length, 0] = 1
        return idx
    lens = rng.integers(min_length, length + 1, size=n_lines)
    pos = np.arange(length + 1)[None, :]
    idx[:, :, 0] = np.where(pos < lens[:, None], idx[:, :, 0], np.where(pos == lens[:, None], 1, -1))
    return idx


def indices_to_lines(idx, i2c):
    """Index arrays [n, T, 1] -> list of strings (up to and including the newline)."""
    cps = np.zeros(len(i2c) + 1, dtype=np.uint32)
    for i, c in i2c.items():
        if len(c) == 1:
            cps[i] = ord(c)
    flat = idx[:, :, 0]
    keep = flat > 0
    counts = keep.sum(axis=1)
    text = cps[flat[keep]].tobytes().decode('utf-32-le')
    ends = np.cumsum(counts)
    return [text[e - c:e] for c, e in zip(counts.tolist(), ends.tolist())]


def layer_shapes(width, depth, voc_size, deep_bidirectional_encoder=False, bridge_dense=False):
    """Keras layer name -> list of weight shapes, in Keras weight order."""
    W, d, V = int(width), int(depth), int(voc_size)
    D = 2 * W if (d == 1 or deep_bidirectional_encoder) else W
    shapes = {'char_input_projection': [(V, W)]}
    for n in range(d):
        bidi = n == 0 or deep_bidirectional_encoder
        in_dim = W if n == 0 else (2 * W if (deep_bidirectional_encoder or n == 1) else W)
        one = [(in_dim, 4 * W), (W, 4 * W), (4 * W,)]
        shapes['encoder_lstm_%d' % (n + 1)] = one + one if bidi else one
        if bridge_dense:
            shapes['bridge_h_%d' % (n + 1)] = [(W, W), (W,)]
            shapes['bridge_c_%d' % (n + 1)] = [(W, W), (W,)]
    shapes['attention_dense'] = [(D, W)]
    for n in range(d - 1):
        shapes['decoder_lstm_%d' % (n + 1)] = [(W, 4 * W), (W, 4 * W), (4 * W,)]
    shapes['decoder_lstm_%d' % d] = [(W, W), (W, 1), (W,), (1,), (W + D, 4 * W), (W, 4 * W), (4 * W,)]
    return shapes


def random_weights(shapes, width, depth, rng, variant="K"):
    """{layer: [arrays]} for `shapes` (layer_shapes) from the initialisers of `variant` (module docstring)."""
    W = int(width)
    noise = 0.1 if variant in ("U", "P", "E") else 0.0

    def glorot(shape):
        lim = np.sqrt(6.0 / (shape[0] + shape[1]))
        return rng.uniform(-lim, lim, size=shape).astype(np.float32)

    def lstm(in_dim):
        q, _ = np.linalg.qr(rng.standard_normal((4 * W, W)))
        bias = np.zeros(4 * W, np.float32)
        bias[W:2 * W] = 1.0
        if noise:
            bias += (rng.standard_normal(4 * W) * noise).astype(np.float32)
        return [glorot((in_dim, 4 * W)), q.T.astype(np.float32), bias]

    out = {}
    for layer, sh in shapes.items():
        if layer == 'char_input_projection':
            E = (rng.standard_normal(sh[0]) * 0.001).astype(np.float32)
            if variant in ("U", "P", "E"):
                E = (E / np.linalg.norm(E, axis=1, keepdims=True)).astype(np.float32)
            if variant in ("P", "E"):
                E = (E * np.float32(12.0)).astype(np.float32)
            if variant == "E":
                E[1] *= np.float32(4.0)
            out[layer] = [E]
        elif layer == 'decoder_lstm_%d' % depth:
            out[layer] = [glorot(sh[0]), glorot(sh[1]), np.zeros(sh[2], np.float32), np.zeros(sh[3], np.float32)] + lstm(sh[4][0])
            if variant == "E":
                out[layer][5] = (out[layer][5] * np.float32(12.0)).astype(np.float32)
        elif 'lstm' in layer:
            out[layer] = sum((lstm(sh[i][0]) for i in range(0, len(sh), 3)), [])
            if variant == "E" and layer.startswith('decoder'):
                out[layer][1] = (out[layer][1] * np.float32(12.0)).astype(np.float32)
        elif len(sh) == 2:      # bridge: Dense kernel + bias
            out[layer] = [glorot(sh[0]), np.zeros(sh[1], np.float32)]
        else:
            out[layer] = [glorot(sh[0])]
    return out

"""Oracle, part 1: topology, seeded weights and the forward pass (numpy).

TEST INFRASTRUCTURE — see ``oracle/__init__.py``.  Parity unpinned.

Restates (file:line under /root/reference/ocrd_cor_asv_ann/lib):
  * ``seq2seq.py:237-314``  encoder graph (input projection, bidirectional bottom
    layer, stacked forward layers, residual rule, bridge, attention init state)
  * ``seq2seq.py:416-480``  one-timestep decoder graph (no residuals, no dropout)
  * ``attention.py:526-575`` DenseAnnotationAttention.attention_call and the
    wiring of ``_call_attend_before`` (:278-305) / ``_get_cell_input`` (:338-344)
  * Keras 2.3.1 ``LSTMCell`` (implementation 2, gate order i,f,c,o, logistic
    recurrent activation forced at seq2seq.py:271) — third-party, not vendored.

All arithmetic is carried out in ``dtype`` (float32 = the reference's precision;
float64 = shadow used by the tests to size tolerances and find near-ties).
"""
from dataclasses import dataclass, field
import numpy as np


@dataclass
class Topology:
    """The persisted / programmatic switches that shape the graphs (seq2seq.py:108-179)."""
    width: int = 512
    depth: int = 2
    voc_size: int = 512
    residual_connections: bool = False
    deep_bidirectional_encoder: bool = False
    bridge_dense: bool = False
    lm_predict: bool = False
    window_width: int = 5
    window_inclusive: bool = True   # K.relu(x, max_value=w, threshold=w)==0  <=>  x <= w

    @property
    def enc_dim(self):
        """D: width of the attended encoder output (seq2seq.py:293-295)."""
        if self.depth == 1 or self.deep_bidirectional_encoder:
            return 2 * self.width
        return self.width

    def layer_is_bidi(self, n):
        return n == 0 or self.deep_bidirectional_encoder

    def enc_in_dim(self, n):
        if n == 0:
            return self.width
        if self.deep_bidirectional_encoder or n == 1:
            return 2 * self.width
        return self.width


# --------------------------------------------------------------------------- weights

def _glorot(rng, shape):
    lim = np.sqrt(6.0 / (shape[0] + shape[1]))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _orthogonal(rng, shape):
    a = rng.standard_normal(shape)
    u, _, vt = np.linalg.svd(a, full_matrices=False)
    q = u if u.shape == shape else vt
    return q.astype(np.float32)


def _lstm_weights(rng, in_dim, W, bias_noise=0.0):
    b = np.zeros(4 * W, np.float32)
    b[W:2 * W] = 1.0          # unit_forget_bias
    if bias_noise:
        # trained models have non-zero biases; with the Keras default (zeros) an LSTM that starts
        # in zero padding stays at exactly zero state, which would hide the un-masked padding
        b += (rng.standard_normal(4 * W) * bias_noise).astype(np.float32)
    return [_glorot(rng, (in_dim, 4 * W)), _orthogonal(rng, (W, 4 * W)), b]


def init_weights(topo, seed=0, variant="U"):
    """Seeded random-init weights in Keras layer / weight order.

    variant "K": Keras default initialisers (embedding N(0, 0.001), seq2seq.py:240).
    variant "U": same, embedding rows rescaled to unit norm (what the regulariser
    at seq2seq.py:549-551 drives trained models to) — gives realistic softmax margins —
    and N(0, 0.1) noise on the LSTM biases (so that padding steps change the state).
    variant "P" ("peaked"): as "U" with embedding rows of norm 12, i.e. logits of standard deviation ~3 like a
    trained model's: the softmax becomes selective (a few candidates above ``beam_threshold_in`` times the
    best instead of none or hundreds), which is what makes the beam search a SEARCH -- tens of live
    hypotheses per line and iteration, finals found among them, early exit by ``beam_width_out``.
    Returns ``{keras_layer_name: [ndarray, ...]}``.
    """
    rng = np.random.default_rng(seed)
    W, d, V, D = topo.width, topo.depth, topo.voc_size, topo.enc_dim
    w = {}
    bn = 0.1 if variant in ("U", "P") else 0.0
    E = (rng.standard_normal((V, W)) * 0.001).astype(np.float32)
    if variant in ("U", "P"):
        E = (E / np.linalg.norm(E, axis=1, keepdims=True)).astype(np.float32)
    if variant == "P":
        E = (E * np.float32(12.0)).astype(np.float32)
    w["char_input_projection"] = [E]
    for n in range(d):
        name = "encoder_lstm_%d" % (n + 1)
        if topo.layer_is_bidi(n):
            w[name] = _lstm_weights(rng, topo.enc_in_dim(n), W, bn) + _lstm_weights(rng, topo.enc_in_dim(n), W, bn)
        else:
            w[name] = _lstm_weights(rng, topo.enc_in_dim(n), W, bn)
        if topo.bridge_dense:
            w["bridge_h_%d" % (n + 1)] = [_glorot(rng, (W, W)), (rng.standard_normal(W) * 0.05).astype(np.float32)]
            w["bridge_c_%d" % (n + 1)] = [_glorot(rng, (W, W)), (rng.standard_normal(W) * 0.05).astype(np.float32)]
    w["attention_dense"] = [_glorot(rng, (D, W))]
    for n in range(d - 1):
        w["decoder_lstm_%d" % (n + 1)] = _lstm_weights(rng, W, W, bn)
    w["decoder_lstm_%d" % d] = [
        _glorot(rng, (W, W)),                                    # W_a
        _glorot(rng, (W, 1)),                                    # v_a
        (rng.standard_normal(W) * 0.05).astype(np.float32),      # b_UW (zeros in Keras; non-zero exercises the path)
        (rng.standard_normal(1) * 0.05).astype(np.float32),      # b_v
    ] + _lstm_weights(rng, W + D, W, bn)
    return w


def cast_weights(w, dtype):
    return {k: [np.asarray(a, dtype) for a in v] for k, v in w.items()}


# --------------------------------------------------------------------------- primitives

def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def lstm_step(x, h, c, K, R, b):
    """Keras LSTMCell.call, implementation 2."""
    W = h.shape[-1]
    z = x @ K + h @ R + b
    i = _sigmoid(z[:, :W])
    f = _sigmoid(z[:, W:2 * W])
    g = np.tanh(z[:, 2 * W:3 * W])
    o = _sigmoid(z[:, 3 * W:])
    c2 = f * c + i * g
    h2 = o * np.tanh(c2)
    return h2, c2


def lstm_scan(x, K, R, b, reverse=False):
    """Full-sequence LSTM over (B,T,in); zero initial state; returns (seq, h_T, c_T).

    ``reverse`` = Keras ``go_backwards`` with the output re-reversed, as
    ``Bidirectional`` does.  No masking: padding steps are ordinary steps
    (SURVEY A.6 item 1).
    """
    B, T, _ = x.shape
    W = R.shape[0]
    h = np.zeros((B, W), x.dtype)
    c = np.zeros((B, W), x.dtype)
    out = np.zeros((B, T, W), x.dtype)
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        h, c = lstm_step(x[:, t], h, c, K, R, b)
        out[:, t] = h
    return out, h, c


def cross_sum(x):
    """seq2seq.py:257-259: add each feature to its neighbour inside adjacent pairs."""
    s = x.shape
    pairs = x.reshape(s[:-1] + (s[-1] // 2, 2))
    return x + pairs[..., ::-1].reshape(s)


def softmax_rows(z):
    z = z - z.max(axis=-1, keepdims=True)
    e = np.exp(z)
    return e / e.sum(axis=-1, keepdims=True)


# --------------------------------------------------------------------------- encoder

def encode(topo, w, x_in):
    """encoder_model.predict_on_batch (seq2seq.py:403-406).

    x_in: (B,T,V) dense input rows (unit vectors, confidences, or all-zero padding).
    Returns ``[enc_out (B,T,D), h_1, c_1, ..., h_d, c_d, a0 (B,T)]``.
    """
    dtype = w["char_input_projection"][0].dtype
    x = np.asarray(x_in, dtype) @ w["char_input_projection"][0]
    states = []
    out = x
    for n in range(topo.depth):
        lw = w["encoder_lstm_%d" % (n + 1)]
        if topo.layer_is_bidi(n):
            inp = out if n == 0 else cross_sum(out)
            fw, _, _ = lstm_scan(inp, *lw[0:3])
            bw, bh, bc = lstm_scan(inp, *lw[3:6], reverse=True)
            out = np.concatenate([fw, bw], axis=-1)
            sh, sc = bh, bc                         # seq2seq.py:280-281
        else:
            out2, sh, sc = lstm_scan(out, *lw[0:3])
            if topo.residual_connections and n >= 2:   # seq2seq.py:284-290 (n==1 skipped)
                out = out2 + out
            else:
                out = out2
        if topo.bridge_dense:
            kh, bh_ = w["bridge_h_%d" % (n + 1)]
            kc, bc_ = w["bridge_c_%d" % (n + 1)]
            sh = np.tanh(sh @ kh + bh_)
            sc = np.tanh(sc @ kc + bc_)
        states += [sh, sc]
    a0 = np.zeros(out.shape[:2], dtype)
    return [out] + states + [a0]


# --------------------------------------------------------------------------- decoder step

def attention(topo, w_att, h_prev, a_prev, enc_out, u):
    """attention.py:526-575, literally over all T positions."""
    W_a, v_a, b_UW, b_v = w_att
    dtype = h_prev.dtype
    T = enc_out.shape[1]
    wq = h_prev @ W_a + b_UW                                        # (B,W)
    e = np.exp(np.tanh(wq[:, None, :] + u) @ v_a + b_v)[..., 0]      # (B,T)
    if topo.window_width > 0:
        steps = np.arange(T, dtype=dtype)
        centre = a_prev @ steps + dtype.type(1)                      # (B,)
        dist = np.abs(centre[:, None] - steps[None, :])
        ww = dtype.type(topo.window_width)
        keep = (dist <= ww) if topo.window_inclusive else (dist < ww)
        e = e * keep.astype(dtype)
    a = e / e.sum(axis=1, keepdims=True)
    ctx = (a[:, :, None] * enc_out).sum(axis=1)
    return ctx, a


def decode_step(topo, w, p_in, enc_out, states, u=None):
    """decoder_model.predict_on_batch for one timestep (seq2seq.py:416-480).

    p_in (B,V): previous softmax row (or zeros at the first step);
    enc_out (B or 1,T,D); states = [h_1,c_1,...,h_d,c_d,a].
    Returns ``(scores (B,V), lm_scores or None, new_states)``.
    ``u`` (= enc_out·U) may be passed in; the reference recomputes it every call (:460).
    """
    E = w["char_input_projection"][0]
    dtype = E.dtype
    d = topo.depth
    if u is None:
        u = enc_out @ w["attention_dense"][0]
    x = np.asarray(p_in, dtype) @ E
    new_states = []
    for n in range(d - 1):
        K, R, b = w["decoder_lstm_%d" % (n + 1)]
        h, c = lstm_step(x, states[2 * n], states[2 * n + 1], K, R, b)
        new_states += [h, c]
        x = h
    top = w["decoder_lstm_%d" % d]
    h_prev, c_prev, a_prev = states[2 * (d - 1)], states[2 * (d - 1) + 1], states[2 * d]
    ctx, a = attention(topo, top[0:4], h_prev, a_prev, enc_out, u)
    K, R, b = top[4:7]
    h, c = lstm_step(np.concatenate([x, ctx], axis=1), h_prev, c_prev, K, R, b)
    scores = softmax_rows(h @ E.T)
    lm_scores = None
    if topo.lm_predict:
        # second evaluation of the top cell with attended == u == 0 (seq2seq.py:464-469):
        # context is exactly zero, so only the x-rows of the cell kernel contribute.
        W = topo.width
        h_lm, _ = lstm_step(x, h_prev, c_prev, K[:W], R, b)
        lm_scores = softmax_rows(h_lm @ E.T)
    new_states += [h, c, a]
    return scores, lm_scores, new_states

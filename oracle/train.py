"""Oracle, part 4: the TRAINING graph, its loss and its gradients (torch float64 autograd on the CPU).

TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.  Parity unpinned (Keras/TF cannot run here): the forward pass is
tied to the rest of the oracle instead -- with dropout off and no residuals, teacher forcing through this graph must
give, step by step, the probabilities ``oracle.model.decode_step`` gives when it is fed the same one-hot history
(``tests/test_train_oracle.py``), and that decoder is the one whose surroundings are pinned to the reference's code.

Restates (file:line under /root/reference/ocrd_cor_asv_ann/lib):
  * ``seq2seq.py:316-390``   training model: teacher-forced decoder over the full target sequence, decoder residual
    adds for n > 0 (:359-360), time-constant dropout between layers (:293-298, :363-367), the attention cell's input
    dropout (:345), tied output projection + softmax (:379)
  * ``seq2seq.py:1061-1117`` teacher forcing arrays: decoder input = target shifted right behind an all-zero start
    vector, output = target, temporal weights 0 on all-zero (padding) output rows
  * ``seq2seq.py:494-497``   loss: Keras ``categorical_crossentropy`` on probabilities (renormalise, clip to
    [1e-7, 1 - 1e-7]) with ``sample_weight_mode='temporal'``: sum(w * ce) / #{w != 0}
  * ``seq2seq.py:530-553``   embedding regulariser (train phase only)
  * ``seq2seq.py:491-497``   ``Adam(clipnorm=5)``: Keras 2.3.1 stand-alone -- every gradient scaled by
    5 / max(5, global norm); lr 1e-3, beta 0.9 / 0.999, epsilon 1e-7, lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t),
    p -= lr_t * m / (sqrt(v) + epsilon)
Not restated: deep bidirectional encoder, bridge, LM loss (``Sequence2Sequence.train`` refuses them).

Dropout masks are INPUTS (already scaled by 1 / keep): ``masks = {"enc": [per encoder layer below the last: (D_n,)],
"dec": [per decoder layer below the top: (W,)], "cell": (B, W + D) or None}``.
"""
import numpy as np
import torch

EPS = 1e-7


def _t(a):
    return torch.tensor(np.asarray(a, np.float64), dtype=torch.float64)


def params_from(w):
    """{layer: [ndarray]} -> {layer: [leaf tensors requiring grad]} (float64)."""
    return {k: [_t(a).requires_grad_(True) for a in v] for k, v in w.items()}


def _lstm_seq(x, K, R, b, h0, c0, reverse=False):
    """x (B,T,in) -> outputs (B,T,W) in ORIGINAL time order, final (h, c) after the last scanned step."""
    B, T, _ = x.shape
    W = R.shape[0]
    h, c = h0, c0
    outs = [None] * T
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        z = x[:, t] @ K + h @ R + b
        i, f, g, o = torch.sigmoid(z[:, :W]), torch.sigmoid(z[:, W:2 * W]), torch.tanh(z[:, 2 * W:3 * W]), torch.sigmoid(z[:, 3 * W:])
        c = f * c + i * g
        h = o * torch.tanh(c)
        outs[t] = h
    return torch.stack(outs, 1), h, c


def forward(topo, p, x_in, dec_in, masks=None):
    """x_in (B,Ts,V), dec_in (B,Td,V) dense arrays (vectorize_lines) -> softmax (B,Td,V) as a torch tensor."""
    assert not topo.deep_bidirectional_encoder and not topo.bridge_dense
    W, d, win = topo.width, topo.depth, topo.window_width
    masks = masks or {}
    x_in, dec_in = _t(x_in), _t(dec_in)
    B, Ts, V = x_in.shape
    E = p["char_input_projection"][0]
    zeros = torch.zeros((B, W), dtype=torch.float64)
    # ---- encoder (seq2seq.py:237-314)
    seq = x_in @ E
    states = []
    for n in range(d):
        ws = p["encoder_lstm_%d" % (n + 1)]
        if n == 0:
            fw, _, _ = _lstm_seq(seq, ws[0], ws[1], ws[2], zeros, zeros)
            bw, hb, cb = _lstm_seq(seq, ws[3], ws[4], ws[5], zeros, zeros, reverse=True)
            out = torch.cat([fw, bw], -1)
            states.append((hb, cb))
        else:
            out, hf, cf = _lstm_seq(seq, ws[0], ws[1], ws[2], zeros, zeros)
            states.append((hf, cf))
            if topo.residual_connections and n >= 2:
                out = out + seq
        if n < d - 1:
            m = (masks.get("enc") or [None] * d)[n]
            if m is not None:
                out = out * _t(m)
        seq = out
    enc_out = seq
    u = enc_out @ p["attention_dense"][0]
    # ---- decoder (seq2seq.py:316-379)
    Td = dec_in.shape[1]
    dec = dec_in @ E
    for n in range(d - 1):
        ws = p["decoder_lstm_%d" % (n + 1)]
        out, _, _ = _lstm_seq(dec, ws[0], ws[1], ws[2], states[n][0], states[n][1])
        if n > 0 and topo.residual_connections:
            out = out + dec
        m = (masks.get("dec") or [None] * d)[n]
        if m is not None:
            out = out * _t(m)
        dec = out
    W_a, v_a, b_UW, b_v, K, R, b = p["decoder_lstm_%d" % d]
    h, c = states[d - 1]
    a = torch.zeros((B, Ts), dtype=torch.float64)
    steps = torch.arange(Ts, dtype=torch.float64)
    cell_mask = masks.get("cell")
    outs = []
    for t in range(Td):
        wq = h @ W_a + b_UW
        e = torch.exp(torch.tanh(wq[:, None, :] + u) @ v_a + b_v)[:, :, 0]
        if win > 0:
            centre = a.detach() @ steps + 1.0          # K.equal / K.cast: no gradient through the window (attention.py:557-566)
            dist = (centre[:, None] - steps[None, :]).abs()
            keep = (dist <= win) if topo.window_inclusive else (dist < win)
            e = e * keep.to(torch.float64)
        a = e / e.sum(1, keepdim=True)
        ctx = (a[:, :, None] * enc_out).sum(1)
        cell_in = torch.cat([dec[:, t], ctx], -1)
        if cell_mask is not None:
            cell_in = cell_in * _t(cell_mask)
        z = cell_in @ K + h @ R + b
        i, f, g, o = torch.sigmoid(z[:, :W]), torch.sigmoid(z[:, W:2 * W]), torch.tanh(z[:, 2 * W:3 * W]), torch.sigmoid(z[:, 3 * W:])
        c = f * c + i * g
        h = o * torch.tanh(c)
        outs.append(h)
    top = torch.stack(outs, 1)
    if d > 1 and topo.residual_connections:
        top = top + dec
    return torch.softmax(top @ E.T, -1)


def regulariser(E):
    """seq2seq.py:530-553 (train phase)."""
    vec0 = E[0]
    mean = E[1:].mean(0).detach()
    norms = (E * E).sum(1)
    return ((vec0 - mean) ** 2).sum() + 0.01 * ((1.0 - norms) ** 2).sum()


def loss_fn(topo, p, x_in, dec_in, dec_out, masks=None, with_regulariser=True):
    """(loss, cross-entropy part) of one batch; dec_out (B,Td,V) one-hot targets with all-zero padding rows."""
    prob = forward(topo, p, x_in, dec_in, masks)
    y = _t(dec_out)
    w = (y.sum(-1) != 0).to(torch.float64)
    q = prob / prob.sum(-1, keepdim=True)
    q = torch.clamp(q, EPS, 1.0 - EPS)
    ce = -(y * torch.log(q)).sum(-1)
    n = torch.clamp(w.sum(), min=1.0)
    xent = (w * ce).sum() / n
    total = xent + (regulariser(p["char_input_projection"][0]) if with_regulariser else 0.0)
    return total, xent


def loss_and_grads(topo, w, x_in, dec_in, dec_out, masks=None, with_regulariser=True):
    """-> (loss, cross-entropy, {layer: [gradient ndarrays]})."""
    p = params_from(w)
    total, xent = loss_fn(topo, p, x_in, dec_in, dec_out, masks, with_regulariser)
    total.backward()
    grads = {k: [(t.grad if t.grad is not None else torch.zeros_like(t)).numpy().copy() for t in v] for k, v in p.items()}
    return float(total), float(xent), grads


def adam_step(w, grads, state, lr=1e-3, b1=0.9, b2=0.999, eps=1e-7, clipnorm=5.0):
    """Keras 2.3.1 ``Adam(clipnorm=5)`` in float64; ``state`` = {"t": int, "m": {...}, "v": {...}} (updated in place).
    Returns the new weights and the global gradient norm."""
    gn = float(np.sqrt(sum(float((g.astype(np.float64) ** 2).sum()) for v in grads.values() for g in v)))
    scale = clipnorm / max(clipnorm, gn) if clipnorm else 1.0
    state["t"] = state.get("t", 0) + 1
    t = state["t"]
    lr_t = lr * np.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
    out = {}
    for k, arrs in w.items():
        out[k] = []
        for i, a in enumerate(arrs):
            g = grads[k][i].astype(np.float64) * scale
            m = state.setdefault("m", {}).setdefault((k, i), np.zeros_like(g))
            v = state.setdefault("v", {}).setdefault((k, i), np.zeros_like(g))
            m[...] = b1 * m + (1.0 - b1) * g
            v[...] = b2 * v + (1.0 - b2) * g * g
            out[k].append((np.asarray(a, np.float64) - lr_t * m / (np.sqrt(v) + eps)).astype(np.float32))
    return out, gn


def teacher_arrays(mapping, voc_size, sources, targets):
    """vectorize_lines (seq2seq.py:1020-1117) for plain strings: dense (B,Ts,V), (B,Td,V), (B,Td,V), weights (B,Td)."""
    c2i = mapping[0]
    B, Ts, Td = len(sources), max(map(len, sources)), max(map(len, targets)) + 1
    x = np.zeros((B, Ts, voc_size), np.uint32)
    di = np.zeros((B, Td, voc_size), np.uint32)
    do = np.zeros((B, Td, voc_size), np.uint32)
    for i, (s, t) in enumerate(zip(sources, targets)):
        for j, ch in enumerate(s):
            x[i, j, c2i.get(ch, 0)] = 1
        for j, ch in enumerate(t):
            idx = c2i.get(ch, 0)
            di[i, j + 1, idx] = 1
            do[i, j, idx] = 1
    w = np.ones(do.shape[:-1], np.float32)
    w[np.all(do == 0, axis=2)] = 0.0
    return x, di, do, w

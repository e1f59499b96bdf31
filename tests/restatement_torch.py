"""Second, independently written restatement of the forward pass (torch, float64).

TEST INFRASTRUCTURE.  Purpose: the reference (Keras 2.3 / TF 1.15) cannot run here, so the numpy
oracle in ``oracle/`` is cross-checked against a restatement that shares no code with it and leans
on a different LSTM implementation: ``torch.nn.LSTM`` / ``torch.nn.LSTMCell`` (gate order i,f,g,o,
the same as Keras; torch keeps (out,in) kernels, Keras (in,out) — /root/reference
ocrd_cor_asv_ann/lib/seq2seq.py:265-302 builds the encoder from Keras LSTM layers).

Follows seq2seq.py:237-314 (encoder), :416-480 (decoder step), attention.py:526-575.
"""
import torch
from torch import nn

DT = torch.float64


def _lstm_module(K, R, b, bidirectional_pair=None):
    """torch.nn.LSTM carrying Keras weights (K (in,4W), R (W,4W), b (4W))."""
    in_dim, W4 = K.shape
    W = W4 // 4
    m = nn.LSTM(in_dim, W, batch_first=True, bidirectional=bidirectional_pair is not None).to(DT)
    with torch.no_grad():
        m.weight_ih_l0.copy_(torch.as_tensor(K, dtype=DT).T)
        m.weight_hh_l0.copy_(torch.as_tensor(R, dtype=DT).T)
        m.bias_ih_l0.copy_(torch.as_tensor(b, dtype=DT))
        m.bias_hh_l0.zero_()
        if bidirectional_pair is not None:
            K2, R2, b2 = bidirectional_pair
            m.weight_ih_l0_reverse.copy_(torch.as_tensor(K2, dtype=DT).T)
            m.weight_hh_l0_reverse.copy_(torch.as_tensor(R2, dtype=DT).T)
            m.bias_ih_l0_reverse.copy_(torch.as_tensor(b2, dtype=DT))
            m.bias_hh_l0_reverse.zero_()
    return m


def _cell_module(K, R, b):
    in_dim, W4 = K.shape
    m = nn.LSTMCell(in_dim, W4 // 4).to(DT)
    with torch.no_grad():
        m.weight_ih.copy_(torch.as_tensor(K, dtype=DT).T)
        m.weight_hh.copy_(torch.as_tensor(R, dtype=DT).T)
        m.bias_ih.copy_(torch.as_tensor(b, dtype=DT))
        m.bias_hh.zero_()
    return m


class Restatement:
    """topo: any object with width/depth/voc_size/residual_connections/deep_bidirectional_encoder/
    bridge_dense/lm_predict/window_width; weights: {keras layer name: [arrays]} (SURVEY A.5 order)."""

    def __init__(self, topo, weights):
        self.t = topo
        self.w = {k: [torch.as_tensor(a, dtype=DT) for a in v] for k, v in weights.items()}
        self.enc = []
        for n in range(topo.depth):
            lw = weights["encoder_lstm_%d" % (n + 1)]
            bidi = n == 0 or topo.deep_bidirectional_encoder
            self.enc.append(_lstm_module(*lw[0:3], bidirectional_pair=lw[3:6] if bidi else None))
        self.dec = [_cell_module(*weights["decoder_lstm_%d" % (n + 1)]) for n in range(topo.depth - 1)]
        self.top = _cell_module(*weights["decoder_lstm_%d" % topo.depth][4:7])

    @torch.no_grad()
    def encode(self, x_dense):
        t = self.t
        seq = torch.as_tensor(x_dense, dtype=DT) @ self.w["char_input_projection"][0]
        states = []
        for n, m in enumerate(self.enc):
            bidi = n == 0 or t.deep_bidirectional_encoder
            if bidi:
                if n > 0:       # seq2seq.py:257-259: neighbour swap inside feature pairs, added
                    sw = seq.reshape(seq.shape[0], seq.shape[1], -1, 2).flip(-1).reshape(seq.shape)
                    seq = seq + sw
                out, (h, c) = m(seq)
                seq = out
                sh, sc = h[1], c[1]                    # backward direction finals
            else:
                out, (h, c) = m(seq)
                seq = out + seq if (t.residual_connections and n >= 2) else out
                sh, sc = h[0], c[0]
            if t.bridge_dense:
                kh, bh = self.w["bridge_h_%d" % (n + 1)]
                kc, bc = self.w["bridge_c_%d" % (n + 1)]
                sh, sc = torch.tanh(sh @ kh + bh), torch.tanh(sc @ kc + bc)
            states += [sh, sc]
        return [seq] + states + [torch.zeros(seq.shape[:2], dtype=DT)]

    @torch.no_grad()
    def step(self, p_in, enc_out, states):
        t, d = self.t, self.t.depth
        E = self.w["char_input_projection"][0]
        W_a, v_a, b_UW, b_v = self.w["decoder_lstm_%d" % d][0:4]
        x = torch.as_tensor(p_in, dtype=DT) @ E
        new = []
        for n, cell in enumerate(self.dec):
            h, c = cell(x, (states[2 * n], states[2 * n + 1]))
            new += [h, c]
            x = h
        h0, c0, a0 = states[2 * d - 2], states[2 * d - 1], states[2 * d]
        T = enc_out.shape[1]
        u = enc_out @ self.w["attention_dense"][0]
        q = h0 @ W_a + b_UW
        energy = torch.exp(torch.einsum("btw,wo->bt", torch.tanh(q[:, None, :] + u), v_a) + b_v)
        if t.window_width > 0:
            pos = torch.arange(T, dtype=DT)
            centre = (a0 * pos).sum(1, keepdim=True) + 1
            energy = torch.where((centre - pos).abs() <= t.window_width, energy, torch.zeros_like(energy))
        a = energy / energy.sum(1, keepdim=True)
        ctx = torch.einsum("bt,btd->bd", a, enc_out.expand(a.shape[0], -1, -1))
        h, c = self.top(torch.cat([x, ctx], 1), (h0, c0))
        scores = torch.softmax(h @ E.T, dim=1)
        lm = None
        if t.lm_predict:
            h_lm, _ = self.top(torch.cat([x, torch.zeros_like(ctx)], 1), (h0, c0))
            lm = torch.softmax(h_lm @ E.T, dim=1)
        return scores, lm, new + [h, c, a]

    @torch.no_grad()
    def greedy_batch(self, x_dense, i2c):
        """decode_batch_greedy (seq2seq.py:1215-1286): strings and per-char probabilities."""
        enc = self.encode(x_dense)
        enc_out, states = enc[0], enc[1:]
        B, T = enc_out.shape[:2]
        live = torch.as_tensor(x_dense).reshape(B, -1).ne(0).any(1)
        p = torch.zeros((B, self.t.voc_size), dtype=DT)
        text, probs = [''] * B, [[] for _ in range(B)]
        for _ in range(2 * T):
            p, _, states = self.step(p, enc_out, states)
            best = p[:, 1:].argmax(1) + 1
            for j in range(B):
                if text[j].endswith('\n') or not live[j]:
                    continue
                text[j] += i2c[int(best[j])]
                probs[j].append(float(p[j, best[j]]))
        return text, probs

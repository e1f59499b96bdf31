"""The training oracle (oracle/train.py, torch float64 autograd) tied to the rest of the oracle: teacher forcing through
the training graph gives the probabilities the one-step decoder graph gives for the same history; its gradients agree
with central differences; its Adam step with a literal transcription of the update rule."""
import numpy as np
import pytest

from oracle import model as om
from oracle import train as ot
from tests.common import make_model, synthetic_lines


def _batch(name, n=4, length=7, seed=5):
    topo, w, mapping = make_model(name)
    src = synthetic_lines(mapping, n, length, seed=seed, min_length=3)
    tgt = synthetic_lines(mapping, n, length, seed=seed + 1, min_length=3)
    return topo, w, mapping, src, tgt


@pytest.mark.parametrize("name", ["c1", "c2s", "c3res"])
def test_teacher_forcing_equals_stepwise_decoder(name):
    topo, w, mapping, src, tgt = _batch(name)
    if topo.bridge_dense:
        topo = om.Topology(width=topo.width, depth=topo.depth, voc_size=topo.voc_size, residual_connections=False)
        w = {k: v for k, v in w.items() if not k.startswith("bridge")}
    x, di, do, wt = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
    w64 = om.cast_weights(w, np.float64)
    prob = ot.forward(topo, ot.params_from(w64), x, di).detach().numpy()
    enc = om.encode(topo, w64, x)
    states = enc[1:]
    for t in range(di.shape[1]):
        scores, _, states = om.decode_step(topo, w64, di[:, t].astype(np.float64), enc[0], states)
        assert np.abs(scores - prob[:, t]).max() < 1e-12, t


def test_gradients_agree_with_central_differences():
    topo, w, mapping, src, tgt = _batch("c2s", n=3, length=5)
    topo = om.Topology(width=16, depth=2, voc_size=topo.voc_size, residual_connections=True)
    w = om.init_weights(topo, 3, "U")
    x, di, do, wt = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
    rng = np.random.default_rng(0)
    masks = {"enc": [rng.choice([0.0, 1.25], size=2 * topo.width), None], "dec": [rng.choice([0.0, 1.25], size=topo.width), None],
             "cell": rng.choice([0.0, 1.25], size=(len(src), 2 * topo.width))}
    loss, xent, _ = ot.loss_and_grads(topo, w, x, di, do, masks)
    assert loss > xent > 0
    # central differences cannot see the regulariser's stop_gradient (seq2seq.py:544): cross-entropy part only here
    _, _, grads = ot.loss_and_grads(topo, w, x, di, do, masks, with_regulariser=False)
    w64 = om.cast_weights(w, np.float64)
    for layer, i in (("char_input_projection", 0), ("encoder_lstm_1", 4), ("encoder_lstm_2", 0), ("attention_dense", 0),
                     ("decoder_lstm_1", 1), ("decoder_lstm_2", 0), ("decoder_lstm_2", 1), ("decoder_lstm_2", 3), ("decoder_lstm_2", 4)):
        a = w64[layer][i]
        for _ in range(3):
            pos = tuple(int(rng.integers(0, s)) for s in a.shape)
            old = a[pos]
            vals = []
            for eps in (1e-5, -1e-5):
                a[pos] = old + eps
                vals.append(float(ot.loss_fn(topo, ot.params_from(w64), x, di, do, masks, with_regulariser=False)[0]))
            a[pos] = old
            num = (vals[0] - vals[1]) / 2e-5
            assert abs(num - grads[layer][i][pos]) < 1e-6 + 1e-4 * abs(num), (layer, i, pos, num, grads[layer][i][pos])


def test_adam_step_literal():
    rng = np.random.default_rng(1)
    w = {"a": [rng.standard_normal((3, 4)).astype(np.float32)], "b": [rng.standard_normal(5).astype(np.float32)]}
    g = {"a": [rng.standard_normal((3, 4)) * 10], "b": [rng.standard_normal(5) * 10]}
    st = {}
    new, gn = ot.adam_step(w, g, st)
    assert gn > 5
    scale = 5.0 / gn
    for k in w:
        gg = g[k][0] * scale
        m, v = 0.1 * gg, 0.001 * gg * gg
        lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
        assert np.allclose(new[k][0], w[k][0] - lr_t * m / (np.sqrt(v) + 1e-7), atol=1e-7)
    new2, _ = ot.adam_step(new, g, st)
    assert st["t"] == 2


def test_regulariser_gradient_stops_at_the_mean():
    """d/dE of (E0 - stopgrad(mean_{i>=1} E_i))^2 + 0.01 sum_i (1 - |E_i|^2)^2 (seq2seq.py:530-553)."""
    import torch
    rng = np.random.default_rng(2)
    E = rng.standard_normal((6, 4))
    t = torch.tensor(E, requires_grad=True)
    ot.regulariser(t).backward()
    want = -0.04 * (1.0 - (E * E).sum(1))[:, None] * E
    want[0] += 2.0 * (E[0] - E[1:].mean(0))
    assert np.allclose(t.grad.numpy(), want, atol=1e-12)

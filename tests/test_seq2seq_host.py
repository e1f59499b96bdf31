"""Host logic of the drop-in class (no GPU): model-file I/O through the built-in HDF5 reader/writer,
CuDNN-layout conversion, vectorisation, batching of gen_lines, error conventions."""
import os

import numpy as np
import pytest

from cor_asv_ann_b200 import h5lite
from cor_asv_ann_b200.seq2seq import Sequence2Sequence, GAP
from oracle import decoders as od
from tests.common import make_model, synthetic_lines


def _s2s_from(topo, w, mapping, status=2):
    s = Sequence2Sequence()
    s.width, s.depth, s.voc_size, s.mapping = topo.width, topo.depth, topo.voc_size, mapping
    s.residual_connections, s.deep_bidirectional_encoder = topo.residual_connections, topo.deep_bidirectional_encoder
    s.bridge_dense, s.lm_predict = topo.bridge_dense, topo.lm_predict
    s.configure()
    for layer, arrs in w.items():
        s.encoder_decoder_model.get_layer(layer).set_weights(arrs)
    s.status = status
    return s


def test_h5lite_roundtrip(tmp_path):
    g = h5lite.Group()
    g.attrs['layer_names'] = np.array([b'a', b'bb_layer'])
    g.attrs['keras_version'] = np.bytes_(b'2.3.1')
    sub = g.require_group('bb_layer/inner')
    sub['kernel:0'] = np.random.default_rng(0).standard_normal((5, 7)).astype(np.float32)
    sub['bias:0'] = np.arange(3, dtype=np.float64)
    g.require_group('config')['flag'] = np.array(True)
    g['config']['n'] = np.array(7)
    g['config']['mapping'] = np.array([0, 10, 97, 228], np.uint32)
    for i in range(40):                      # more entries than one h5py-default symbol node would hold
        g.require_group('many')['d%02d' % i] = np.full((2,), i, np.int32)
    path = str(tmp_path / 't.h5')
    h5lite.write(path, g)
    r = h5lite.read(path)
    assert list(r.attrs['layer_names']) == [b'a', b'bb_layer']
    assert np.array_equal(r['bb_layer']['inner']['kernel:0'], sub['kernel:0'])
    assert r['bb_layer']['inner']['bias:0'].dtype == np.float64
    assert r['config']['flag'][()] == True and r['config']['flag'].dtype == bool   # noqa: E712
    assert int(r['config']['n'][()]) == 7
    assert list(r['config']['mapping']) == [0, 10, 97, 228]
    assert sorted(r['many']) == ['d%02d' % i for i in range(40)] and r['many']['d17'][0] == 17
    with open(path, 'rb') as f:
        assert f.read(8) == b'\x89HDF\r\n\x1a\n'


@pytest.mark.parametrize("name", ["c1", "c3res", "c4s"])
def test_save_load_model_file(name, tmp_path):
    topo, w, mapping = make_model(name)
    s = _s2s_from(topo, w, mapping)
    path = str(tmp_path / 'model.h5')
    s.save(path)
    t = Sequence2Sequence()
    t.load_config(path)
    assert (t.width, t.depth, t.voc_size) == (topo.width, topo.depth, topo.voc_size)
    assert t.residual_connections == topo.residual_connections and t.bridge_dense == topo.bridge_dense
    assert t.mapping[0] == mapping[0] and t.mapping[1] == mapping[1]
    t.configure()
    assert t.status == 1
    t.load_weights(path)
    assert t.status == 2
    for layer, arrs in w.items():
        for a, b in zip(arrs, t._weights[layer]):
            assert np.array_equal(np.asarray(a, np.float32), b)
    root = h5lite.read(path)     # Keras save_weights layout
    assert [n.decode() for n in root.attrs['layer_names']][0] == 'char_input_projection'
    top = 'decoder_lstm_%d' % topo.depth
    assert [n.decode().split('/')[-1] for n in root[top].attrs['weight_names']][:4] == ['W_a:0', 'v_a:0', 'b_UW:0', 'b_v:0']


def test_cudnn_layout_is_converted_on_load(tmp_path):
    """A GPU-trained layer is stored in CuDNNLSTM layout (bias 8W, per-gate transposed kernels)."""
    topo, w, mapping = make_model("c2s")
    s = _s2s_from(topo, w, mapping)
    W = topo.width

    def to_cudnn(k, r, b):          # Keras _convert_rnn_weights, LSTM -> CuDNNLSTM
        k2 = np.hstack([g.T.reshape(g.shape, order='C') for g in np.hsplit(k, 4)])
        r2 = np.hstack([g.T for g in np.hsplit(r, 4)])
        return [k2, r2, np.concatenate([b * 0.25, b * 0.75])]

    orig = [np.array(a) for a in s._weights['encoder_lstm_2']]
    s._weights['encoder_lstm_2'] = to_cudnn(*orig)
    bw = [np.array(a) for a in s._weights['encoder_lstm_1']]
    s._weights['encoder_lstm_1'] = bw[:3] + to_cudnn(*bw[3:])
    path = str(tmp_path / 'cudnn.h5')
    s.save(path)
    t = Sequence2Sequence()
    t.load_config(path)
    t.configure()
    t.load_weights(path)
    for a, b in zip(orig, t._weights['encoder_lstm_2']):
        assert np.allclose(a, b, atol=1e-7)
    for a, b in zip(bw, t._weights['encoder_lstm_1']):
        assert np.allclose(a, b, atol=1e-7)


def test_load_weights_shape_mismatch_raises(tmp_path):
    topo, w, mapping = make_model("c1")
    s = _s2s_from(topo, w, mapping)
    path = str(tmp_path / 'm.h5')
    s.save(path)
    t = Sequence2Sequence()
    t.load_config(path)
    t.width = 64
    t.configure()
    with pytest.raises(ValueError):
        t.load_weights(path)


def test_status_assertions_and_empty_input():
    s = Sequence2Sequence()
    with pytest.raises(AssertionError):
        s.save('/tmp/never.h5')
    with pytest.raises(AssertionError):
        s.load_weights('/tmp/never.h5')
    with pytest.raises(AssertionError):
        next(s.predict([]))
    assert s.correct_lines([]) == ([], [], [], [])
    with pytest.raises(AssertionError):
        s.correct_lines(['a\n'], fast=True, greedy=False)
    assert 'status: empty' in repr(s)


def test_vectorize_lines_matches_oracle():
    topo, w, mapping = make_model("c1")
    s = _s2s_from(topo, w, mapping)
    lines = synthetic_lines(mapping, 4, 9, seed=3, min_length=2) + ['\x01unmapped\n']
    enc, dec_in, dec_out, weights = s.vectorize_lines(lines, lines)
    assert enc.dtype == np.uint32 and (enc == od.vectorize_lines(mapping, topo.voc_size, lines)).all()
    assert dec_in.shape == (5, enc.shape[1] + 1, topo.voc_size)
    assert (dec_in[:, 1:] == dec_out[:, :-1]).all() and weights[0].sum() == len(lines[0])
    conf = [[0.5 + 0.01 * j for j, _ in enumerate(l)] for l in lines]
    enc2 = s.vectorize_lines(lines, lines, conf)[0]
    assert enc2.dtype == np.float32 and np.array_equal(enc2, od.vectorize_lines(mapping, topo.voc_size, lines, conf))


def test_gen_lines_batches_in_file_order_and_pads(tmp_path):
    s = Sequence2Sequence()
    s.batch_size = 4
    p1, p2 = tmp_path / 'a.txt', tmp_path / 'b.txt'
    p1.write_text('eins\tEins\nzwei\tZwei\ndrei\n')
    p2.write_text('vier\nfünf\nsechs\n')
    batches = list(s.gen_lines([str(p1), str(p2)], repeat=False, unsupervised=True))
    assert len(batches) == 2
    src, conf, tgt, names = batches[0]
    assert src == ['eins\n', 'zwei\n', 'drei\n', 'vier\n'] and conf is None and tgt == src
    assert names == [str(p1)] * 3 + [str(p2)]
    src, _, _, names = batches[1]
    assert src == ['fünf\n', 'sechs\n', '', ''] and names[2:] == [None, None]     # NFC, '' padding

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


# ---------------------------------------------------------------------------------------------- foreign HDF5 bytes
# Files assembled by tests/h5_foreign.py: a second, independent encoder that follows the HDF5 File Format
# Specification and makes libhdf5's layout choices (multi-leaf and two-level group B-trees at LEAF_K = 4, object
# header continuation blocks and NIL messages, modification-time messages, cached symbol table entries, dead heap
# names, maximum dimensions, compact and big-endian datasets, |S20 and variable-length string attributes, holes).
# Not a byte written by h5lite's own writer goes through its reader here.  The reference's shipped models
# (wrapper/ocrd-tool.json:66,72) live in an absent submodule and stay unverified.
from tests import h5_foreign as hf


def _foreign_tree():
    rng = np.random.default_rng(3)
    many = {"d%02d" % i: hf.Dset(np.full((3,), i, np.int32), maxdims=bool(i % 2)) for i in range(45)}     # 9 leaves -> level-1 root
    root = hf.Grp(
        children={
            "many": hf.Grp(children=many, dead_names=("deleted_link", "another_deleted")),
            "few": hf.Grp(children={"kernel:0": hf.Dset(rng.standard_normal((5, 7)).astype(np.float32), maxdims=True),
                                    "be": hf.Dset(np.arange(6, dtype=np.float64).reshape(2, 3), big_endian=True),
                                    "tiny": hf.Dset(np.array([1, 2, 3], np.int16), layout="compact"),
                                    "empty": hf.Dset(np.zeros((0, 4), np.float32))},
                          attrs={"weight_names": np.array([b"kernel:0", b"be"]), "note": hf.VlenStr("variable-length")}),
            "config": hf.Grp(children={"width": hf.Dset(np.array(320)), "stateful": hf.Dset(np.array(False)),
                                       "flag": hf.Dset(np.array(True), layout="compact"),
                                       "mapping": hf.Dset(np.array([0, 10, 97, 228, 8364], np.uint32))}),
        },
        attrs={"layer_names": np.array([b"few", b"many"]).astype("S20"), "keras_version": np.bytes_(b"2.3.1"),
               "backend": np.bytes_(b"tensorflow"), "a1": np.arange(4, dtype=np.int64), "a2": np.float32(2.5),
               "a3": np.array([True, False]), "a4": np.array([1.5, -2.5])})
    return root, many, rng


@pytest.mark.parametrize("kw", [dict(), dict(hole=0, first_block=8, split_at=32), dict(hole=40, first_block=1, split_at=2)])
def test_h5lite_reads_libhdf5_style_files(tmp_path, kw):
    root, many, _ = _foreign_tree()
    path = str(tmp_path / "foreign.h5")
    hf.write(path, root, **kw)
    r = h5lite.read(path)
    assert sorted(r) == ["config", "few", "many"]
    assert sorted(r["many"]) == sorted(many)                   # every leaf of the (two-level) B-tree, no dead names
    for name, d in many.items():
        assert np.array_equal(r["many"][name], d.array) and r["many"][name].dtype == np.int32
    few = root.children["few"].children
    assert np.array_equal(r["few"]["kernel:0"], few["kernel:0"].array) and r["few"]["kernel:0"].dtype == np.float32
    assert np.array_equal(r["few"]["be"], few["be"].array) and r["few"]["be"].dtype.itemsize == 8      # big-endian on disk
    assert list(r["few"]["tiny"]) == [1, 2, 3]
    assert r["few"]["empty"].shape == (0, 4)
    assert [bytes(x) for x in r["few"].attrs["weight_names"]] == [b"kernel:0", b"be"]
    assert "note" in r["few"].attrs and r["few"].attrs["note"] is None       # variable-length string: skipped, not fatal
    assert int(r["config"]["width"][()]) == 320 and r["config"]["width"].shape == ()
    assert r["config"]["stateful"][()] == False and r["config"]["stateful"].dtype == bool      # noqa: E712
    assert r["config"]["flag"][()] == True                                                      # noqa: E712
    assert list(r["config"]["mapping"]) == [0, 10, 97, 228, 8364]
    assert [bytes(x) for x in r.attrs["layer_names"]] == [b"few", b"many"]    # |S20: numpy strips the null padding
    assert r.attrs["layer_names"].dtype == np.dtype("S20")
    assert bytes(r.attrs["keras_version"]) == b"2.3.1"
    assert list(r.attrs["a1"]) == [0, 1, 2, 3] and float(r.attrs["a2"]) == 2.5
    assert list(r.attrs["a3"]) == [True, False] and list(r.attrs["a4"]) == [1.5, -2.5]


def _keras_names(s, layer):
    return s._weight_names(layer, len(s._weights[layer]), s.depth)


@pytest.mark.parametrize("name,nest", [("c1", None), ("c3res", None), ("c4s", "model_weights")])
def test_load_model_from_libhdf5_style_file(name, nest, tmp_path):
    """load_config / load_weights on a model file laid out like Keras 2.3.1 + h5py 2.x write it (save_weights layout,
    or nested under 'model_weights' like Model.save), assembled by the independent encoder."""
    topo, w, mapping = make_model(name)
    s = _s2s_from(topo, w, mapping)
    config = {"width": topo.width, "depth": topo.depth, "stateful": False,
              "residual_connections": topo.residual_connections, "deep_bidirectional_encoder": topo.deep_bidirectional_encoder,
              "bridge_dense": topo.bridge_dense,
              "mapping": np.fromiter((ord(mapping[1][i]) if mapping[1][i] else 0 for i in range(topo.voc_size)), dtype=np.uint32)}
    root = hf.keras_weight_file({k: s._weights[k] for k in s._layer_shapes()}, {k: _keras_names(s, k) for k in s._layer_shapes()},
                                config=None if nest else config, nest=nest)
    if nest:         # Model.save nests the weights; the reference's config group sits at the top level (seq2seq.py:1130)
        root.children["config"] = hf.Grp(children={k: hf.Dset(np.asarray(v)) for k, v in config.items()})
    path = str(tmp_path / "keras_model.h5")
    hf.write(path, root)
    t = Sequence2Sequence()
    t.load_config(path)
    assert (t.width, t.depth, t.voc_size) == (topo.width, topo.depth, topo.voc_size)
    assert t.residual_connections == topo.residual_connections and t.deep_bidirectional_encoder == topo.deep_bidirectional_encoder
    assert t.mapping[1] == mapping[1]
    t.lm_predict = topo.lm_predict
    t.configure()
    t.load_weights(path)
    assert t.status == 2
    for layer, arrs in w.items():
        for a, b in zip(arrs, t._weights[layer]):
            assert np.array_equal(np.asarray(a, np.float32), b), layer


def test_transfer_from_renamed_lm_file(tmp_path):
    """transfer-dta-lm.py:1-42: an ocrd_keraslm weight file whose lstm_1 / lstm_2 groups were COPIED to decoder_lstm_1 /
    decoder_lstm_2 (so the datasets still sit under the old inner names and weight_names still says lstm_1/kernel:0),
    whose first kernel lost its 10 context rows, and whose layer_names attribute was rewritten as |S20.
    load_transfer_weights takes the lower decoder layer and skips what does not fit (skip_mismatch=True)."""
    topo, w, mapping = make_model("c2s")
    W, V = topo.width, topo.voc_size
    rng = np.random.default_rng(5)
    lm = {"decoder_lstm_1": [rng.standard_normal((W, 4 * W)).astype(np.float32), rng.standard_normal((W, 4 * W)).astype(np.float32),
                             rng.standard_normal(4 * W).astype(np.float32)],
          "decoder_lstm_2": [rng.standard_normal((W, 4 * W)).astype(np.float32), rng.standard_normal((W, 4 * W)).astype(np.float32),
                             rng.standard_normal(4 * W).astype(np.float32)],
          "dense_1": [rng.standard_normal((W, V)).astype(np.float32), rng.standard_normal(V).astype(np.float32)]}
    names = {"decoder_lstm_1": ["lstm_1/kernel:0", "lstm_1/recurrent_kernel:0", "lstm_1/bias:0"],
             "decoder_lstm_2": ["lstm_2/kernel:0", "lstm_2/recurrent_kernel:0", "lstm_2/bias:0"],
             "dense_1": ["dense_1/kernel:0", "dense_1/bias:0"]}
    root = hf.keras_weight_file(lm, names, layer_name_width=20)
    for g, dead in (("decoder_lstm_1", "lstm_1"), ("decoder_lstm_2", "lstm_2")):
        root.children[g].children[dead].dead_names = ("kernel:0",)      # `del f[...]; create_dataset(...)` leaves the old name in the heap
    root.dead_names = ("lstm_1", "lstm_2")
    path = str(tmp_path / "lm-char.h5")
    hf.write(path, root, hole=56)
    s = _s2s_from(topo, w, mapping, status=1)
    before = {k: [np.array(a) for a in v] for k, v in s._weights.items()}
    s.load_transfer_weights(path)
    assert s.status == 1
    for a, b in zip(lm["decoder_lstm_1"], s._weights["decoder_lstm_1"]):
        assert np.array_equal(a, b)
    # the LM's top layer has no context rows (W x 4W against (W + D) x 4W): skipped, like every layer the file lacks
    for layer in ("decoder_lstm_2", "encoder_lstm_1", "char_input_projection"):
        for a, b in zip(before[layer], s._weights[layer]):
            assert np.array_equal(a, b), layer

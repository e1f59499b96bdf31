#!/usr/bin/env python
"""Run the REFERENCE's own Python code for the hot path in this container and freeze its outputs.

TEST INFRASTRUCTURE.  Needs /root/reference (build container only); the GPU box sees only the
committed fixtures ``tests/golden/ref_*.npz`` this script writes.

What can and cannot be executed here
------------------------------------
The reference's neural arithmetic is Keras 2.3 / TensorFlow 1.15 (not installable), but
everything AROUND ``predict_on_batch`` is plain Python + numpy and runs unmodified once the two
missing imports of ``lib/seq2seq.py`` (``h5py``, and ``uniseg`` via ``lib/alignment.py``) are
satisfied by empty stub modules:

  * ``Sequence2Sequence.correct_lines / vectorize_lines / decode_batch_greedy /
    decode_sequence_greedy / decode_sequence_beam`` and ``Node``  (lib/seq2seq.py:782-842,
    1020-1119, 1215-1608) — loaded from the reference file and run as is; the two Keras models
    ``encoder_model`` / ``decoder_model`` are replaced by objects whose ``predict_on_batch`` calls the
    oracle's numpy network (oracle/model.py) with the reference's exact input/output lists
    (seq2seq.py:403-406, 477-480).
  * ``DenseAnnotationAttention.attention_call`` (lib/attention.py:526-575) — loaded from the
    reference file and run as is on a numpy implementation of the dozen ``keras.backend`` functions
    it calls (``K.relu(max_value, threshold)`` follows Keras 2.3.1 ``tensorflow_backend.relu``,
    SURVEY.md A.4).
  * ``_alignment2path`` (wrapper/transcode.py:279-349) — the function's source is cut out of the
    reference file with ``ast`` (the module itself imports ``ocrd``) and executed as is.

So the decoders, the search, the vectorisation, the attention formula and the Viterbi re-alignment
are pinned against the reference's own code; what stays a restatement is the Keras LSTM cell /
Dense / softmax arithmetic inside ``predict_on_batch``.

numpy: the reference pins numpy<1.24 (legacy scalar promotion); this image has numpy 2.x (NEP 50).
The reference code run here therefore accumulates beam costs in float32 where its own environment
used float64.  The oracle has a switch for exactly those lines (``oracle.decoders.PROMOTION``);
the comparison with the reference is made under "nep50", the product follows "legacy".
"""
import ast
import importlib.util
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
REF = "/root/reference/ocrd_cor_asv_ann"
GOLDEN = os.path.join(ROOT, "tests", "golden")


def available():
    return os.path.isfile(os.path.join(REF, "lib", "seq2seq.py"))


# ----------------------------------------------------------------------------- loading

_MODS = {}


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load_lib():
    """lib/alignment.py and lib/seq2seq.py of the reference, imported from where they lie."""
    if "seq2seq" in _MODS:
        return _MODS["seq2seq"]
    sys.dont_write_bytecode = True
    if "h5py" not in sys.modules:
        _stub("h5py")
    if "uniseg" not in sys.modules:
        wb = _stub("uniseg.wordbreak")
        _stub("uniseg", wordbreak=wb)
    pkg = types.ModuleType("cor_ref_lib")
    pkg.__path__ = [os.path.join(REF, "lib")]
    sys.modules["cor_ref_lib"] = pkg
    for name in ("alignment", "seq2seq"):
        spec = importlib.util.spec_from_file_location("cor_ref_lib." + name, os.path.join(REF, "lib", name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["cor_ref_lib." + name] = mod
        spec.loader.exec_module(mod)
        _MODS[name] = mod
    return _MODS["seq2seq"]


def load_alignment2path():
    """wrapper/transcode.py:_alignment2path, cut out of the file (the module imports ocrd)."""
    if "a2p" in _MODS:
        return _MODS["a2p"]
    path = os.path.join(REF, "wrapper", "transcode.py")
    src = open(path, encoding="utf-8").read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "_alignment2path")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    _MODS["a2p"] = ns["_alignment2path"]
    return _MODS["a2p"]


class _NumpyBackend:
    """The keras.backend functions attention_call uses (attention.py:539-575), on numpy float32."""

    @staticmethod
    def dot(x, y):
        return np.dot(x, y)

    @staticmethod
    def repeat(x, n):                       # (B,W) -> (B,n,W)
        return np.repeat(x[:, None, :], int(n), axis=1)

    @staticmethod
    def shape(x):
        return x.shape

    exp = staticmethod(np.exp)
    tanh = staticmethod(np.tanh)
    abs = staticmethod(np.abs)
    zeros_like = staticmethod(np.zeros_like)
    equal = staticmethod(np.equal)

    @staticmethod
    def arange(n, dtype='int32'):
        return np.arange(int(n), dtype=dtype)

    @staticmethod
    def expand_dims(x, axis=-1):
        return np.expand_dims(x, axis)

    @staticmethod
    def squeeze(x, axis):
        return np.squeeze(x, axis)

    @staticmethod
    def cast(x, dtype):
        return x.astype(dtype)

    @staticmethod
    def dtype(x):
        return x.dtype.name

    @staticmethod
    def sum(x, axis=None, keepdims=False):
        return np.sum(x, axis=axis, keepdims=keepdims)

    @staticmethod
    def relu(x, alpha=0., max_value=None, threshold=0.):
        # Keras 2.3.1 tensorflow_backend.relu with alpha == 0 and threshold != 0:
        #   x = x * cast(greater(x, threshold), floatx); then clip_by_value(x, 0, max_value)
        assert alpha == 0.
        if threshold != 0:
            x = x * (x > threshold).astype(x.dtype)
        else:
            x = np.maximum(x, 0)
        if max_value is not None:
            x = np.clip(x, 0, max_value)
        return x


def load_attention_call():
    """lib/attention.py:DenseAnnotationAttention.attention_call, with keras stubbed out."""
    if "attn" in _MODS:
        return _MODS["attn"]
    sys.dont_write_bytecode = True
    saved = {k: sys.modules.get(k) for k in list(sys.modules) if k == "keras" or k.startswith("keras.")}

    class _Anything:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, name):
            return _Anything()

        def __call__(self, *a, **k):
            return _Anything()

    K = _NumpyBackend()
    keras = _stub("keras", backend=K, initializers=_Anything(), regularizers=_Anything(), constraints=_Anything())
    _stub("keras.backend")
    sys.modules["keras.backend"] = K
    _stub("keras.engine")
    _stub("keras.engine.base_layer", _collect_previous_mask=lambda *a: None, disable_tracking=lambda f: f)
    _stub("keras.layers", Layer=type("Layer", (), {"__init__": lambda self, *a, **k: None}),
          InputSpec=_Anything, concatenate=lambda *a, **k: None)
    _stub("keras.utils")
    _stub("keras.utils.generic_utils", has_arg=lambda *a: False, to_list=lambda x: list(x))
    keras.layers = sys.modules["keras.layers"]
    try:
        spec = importlib.util.spec_from_file_location("cor_ref_attention", os.path.join(REF, "lib", "attention.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k in [k for k in sys.modules if k == "keras" or k.startswith("keras.")]:
            del sys.modules[k]
        for k, v in saved.items():
            if v is not None:
                sys.modules[k] = v
    _MODS["attn"] = mod.DenseAnnotationAttention.attention_call
    return _MODS["attn"]


# ----------------------------------------------------------------------------- the stand-in models

class _EncoderModel:
    """encoder_model (seq2seq.py:403-406): x (B,T,V) -> [enc_out, h1, c1, ..., hd, cd, a0]."""

    def __init__(self, topo, w):
        self.topo, self.w = topo, w

    def predict_on_batch(self, x):
        from oracle.model import encode
        return encode(self.topo, self.w, x)


class _DecoderModel:
    """decoder_model (seq2seq.py:477-480): [p (B,1,V), enc_out] + states -> [scores (B,1,V), (lm)] + states."""

    def __init__(self, topo, w):
        self.topo, self.w = topo, w

    def predict_on_batch(self, inputs):
        from oracle.model import decode_step
        p, enc_out = inputs[0], inputs[1]
        states = list(inputs[2:])
        scores, lm, new_states = decode_step(self.topo, self.w, np.asarray(p)[:, -1], enc_out, states)
        out = [scores[:, None, :]]
        if lm is not None:
            out.append(lm[:, None, :])
        return out + list(new_states)


def reference_model(topo, w, mapping, **params):
    """A reference ``Sequence2Sequence`` in status 2 whose two Keras models are the stand-ins above."""
    import logging
    mod = load_lib()
    log = logging.getLogger("cor_ref")
    log.setLevel(logging.CRITICAL)
    s2s = mod.Sequence2Sequence(logger=log, progbars=False)
    s2s.width, s2s.depth, s2s.voc_size = topo.width, topo.depth, topo.voc_size
    s2s.mapping = mapping
    s2s.residual_connections = topo.residual_connections
    s2s.deep_bidirectional_encoder = topo.deep_bidirectional_encoder
    s2s.bridge_dense = topo.bridge_dense
    s2s.lm_predict = topo.lm_predict
    for k, v in params.items():
        assert hasattr(s2s, k), k
        setattr(s2s, k, v)
    s2s.encoder_model = _EncoderModel(topo, w)
    s2s.decoder_model = _DecoderModel(topo, w)
    s2s.status = 2
    return s2s


# ----------------------------------------------------------------------------- fixture cases

# name -> (config, #lines, max length, min length, seed, mode, beam/rejection parameters)
DECODER_CASES = {
    "fast_c1": ("c1", 12, 30, 9, 311, "fast", {}),
    "fast_c2s": ("c2s", 10, 24, 6, 312, "fast", {}),
    "greedy_c2s": ("c2s", 6, 20, 8, 313, "greedy", {}),
    "greedy_c1": ("c1", 5, 18, 7, 314, "greedy", {}),
    "fast_c3res": ("c3res", 5, 18, 7, 314, "fast", {}),   # per-line greedy on this model hits the reference's NaN-feedback crash
    "beam_c1_default": ("c1", 4, 14, 8, 315, "beam", dict(rejection_threshold=0.3)),
    "beam_c2s_rej05_w4": ("c2s", 4, 12, 7, 316, "beam", dict(rejection_threshold=0.5, beam_width_in=4)),
    "beam_c2s_rej0": ("c2s", 3, 8, 5, 317, "beam", dict(rejection_threshold=0, beam_width_in=3, batch_size=16)),
    "beam_c2s_rej01": ("c2s", 3, 10, 6, 318, "beam", dict(rejection_threshold=0.1, beam_width_in=5, batch_size=32)),
    "beam_c4s_lm": ("c4s", 3, 9, 6, 319, "beam", dict(rejection_threshold=0.3, beam_width_in=4, batch_size=32)),
}


# confidence / confusion-network input (vectorize_lines :1041-1092): glyph alternatives of the
# reference's own test pages (tests/golden/kant_lines.json), config 1 over their alphabet
KANT_CASES = {
    "kant_confmat_fast": (0, slice(0, 8), "fast", {}),
    "kant_confmat_greedy": (1, slice(3, 7), "greedy", {}),
    "kant_confmat_beam": (0, slice(8, 11), "beam", dict(rejection_threshold=0.3, beam_width_in=4, batch_size=32)),
}


def case_inputs(name):
    from tests.common import make_model, synthetic_lines, kant_fixture, kant_model
    if name in KANT_CASES:
        pno, sl, mode, params = KANT_CASES[name]
        alphabet, pages = kant_fixture()
        topo, w, mapping = kant_model(alphabet)
        lines = [[[(s, c) for s, c in chunk] for chunk in line] for line in pages[pno][sl]]
        return topo, w, mapping, lines, mode, params
    cfg, n, length, min_length, seed, mode, params = DECODER_CASES[name]
    topo, w, mapping = make_model(cfg)
    lines = synthetic_lines(mapping, n, length, seed=seed, min_length=min_length)
    return topo, w, mapping, lines, mode, params


def run_reference_case(name, conf=None, lines=None):
    """correct_lines of the REFERENCE class on the case's lines -> (texts, probs, scores, alignments)."""
    topo, w, mapping, lines0, mode, params = case_inputs(name)
    s2s = reference_model(topo, w, mapping, **params)
    lines = lines0 if lines is None else lines
    if name in KANT_CASES:
        conf = lines                                  # the wrapper passes the confmats twice (transcode.py:112)
    fast = mode == "fast"
    greedy = mode != "beam"
    with np.errstate(divide='ignore', invalid='ignore'):
        return s2s.correct_lines(lines, conf=conf, fast=fast, greedy=greedy)


def run_oracle_case(name, conf=None, lines=None, promotion="nep50"):
    from oracle import decoders as od
    topo, w, mapping, lines0, mode, params = case_inputs(name)
    lines = lines0 if lines is None else lines
    if name in KANT_CASES:
        conf = lines
    old = od.PROMOTION
    od.PROMOTION = promotion
    try:
        return od.correct_lines(topo, w, mapping, lines, conf=conf, fast=(mode == "fast"),
                                greedy=(mode != "beam"), beam_params=params if mode == "beam" else None)
    finally:
        od.PROMOTION = old


def pack_result(res):
    """correct_lines 4-tuple -> flat arrays for an .npz."""
    texts, probs, scores, aligns = res
    T = max((len(a[0]) for a in aligns if len(a)), default=0)
    return {
        "texts": np.array([t.encode("utf-8") for t in texts], dtype=object).astype("S"),
        "probs": np.array([float(p) for pl in probs for p in pl], np.float64),
        "lens": np.array([len(pl) for pl in probs], np.int64),
        "scores": np.array([float(s) for s in scores], np.float64),
        "align_centre": np.array([float(np.dot(np.asarray(v, np.float64), np.arange(len(v))))
                                  for al in aligns for v in al], np.float64),
        "align_T": np.array([T], np.int64),
    }


def viterbi_cases():
    """(alignment list, i_max, j_max, min_score) inputs for _alignment2path: soft attention rows
    from the decoders above plus synthetic band matrices that hit the i_max <> j_max branches."""
    rng = np.random.default_rng(991)
    cases = []
    for (S, T, i_max, j_max) in [(12, 12, 12, 12), (20, 16, 14, 18), (15, 22, 21, 13), (9, 9, 8, 9),
                                 (30, 30, 30, 30), (6, 11, 11, 5), (11, 6, 5, 11), (3, 3, 3, 3), (2, 5, 4, 2)]:
        a = np.zeros((S, T), np.float32)
        for j in range(S):
            c = min(T - 1, max(0, int(round(j * (T - 1) / max(S - 1, 1) + rng.normal(0, 0.7)))))
            lo, hi = max(0, c - 2), min(T, c + 3)
            v = rng.random(hi - lo).astype(np.float32) ** 2
            a[j, lo:hi] = v / v.sum()
        cases.append((a, i_max, j_max, np.float32(1.0 / 80)))
    # hard one-hot alignments (rejection path) and ties
    eye = np.eye(10, dtype=np.float32)
    cases.append((eye, 10, 10, np.float32(1.0 / 80)))
    cases.append((np.full((7, 7), 1.0 / 7, np.float32), 7, 7, np.float32(1.0 / 80)))
    return cases


def run_reference_viterbi(a, i_max, j_max, min_score):
    fn = load_alignment2path()
    realignment, dist = fn([row for row in a], i_max, j_max, min_score)
    out = np.full(i_max + 1, -1, np.int64)
    extra = {}
    for k, v in realignment.items():
        if 0 <= k <= i_max:
            out[k] = v
        else:
            extra[k] = v
    return out, float(dist), extra


def attention_case(seed, B=5, T=17, W=12, D=20, after_rejection=False):
    rng = np.random.default_rng(seed)
    f = lambda *s: rng.normal(0, 0.5, s).astype(np.float32)
    h = f(B, W)
    enc = f(B, T, D)
    u = f(B, T, W)
    a_prev = np.zeros((B, T), np.float32)
    for b in range(B):
        if after_rejection or b == 0:
            a_prev[b, rng.integers(0, T)] = 1.0                   # integer centre: the inclusive edge of the window
        elif b == 1:
            pass                                                  # first step: a_prev == 0
        else:
            c = rng.integers(0, T - 3)
            v = rng.random(3).astype(np.float32)
            a_prev[b, c:c + 3] = v / v.sum()
    return dict(h=h, enc=enc, u=u, a_prev=a_prev, W_a=f(W, W), v_a=f(W, 1), b_UW=f(W), b_v=f(1))


def run_reference_attention(c, window_width=5):
    call = load_attention_call()
    me = types.SimpleNamespace(W_a=c["W_a"], v_a=c["v_a"], b_UW=c["b_UW"], b_v=c["b_v"], window_width=window_width)
    ctx, (a,) = call(me, None, [c["h"], None], [c["enc"], c["u"]], [c["a_prev"]], [None])
    return np.asarray(ctx), np.asarray(a)


# ----------------------------------------------------------------------------- fixture writer

def main():
    assert available(), "needs /root/reference"
    os.makedirs(GOLDEN, exist_ok=True)
    for name in list(DECODER_CASES) + list(KANT_CASES):
        res = run_reference_case(name)
        np.savez_compressed(os.path.join(GOLDEN, "ref_%s.npz" % name), **pack_result(res))
        print(name, [t.strip() for t in res[0]][:3], [round(float(s), 4) for s in res[2]][:3])
    vit = {}
    for k, (a, i_max, j_max, ms) in enumerate(viterbi_cases()):
        path, dist, extra = run_reference_viterbi(a, i_max, j_max, ms)
        vit["a%d" % k] = a
        vit["dims%d" % k] = np.array([i_max, j_max], np.int64)
        vit["ms%d" % k] = np.array([ms], np.float32)
        vit["path%d" % k] = path
        vit["dist%d" % k] = np.array([dist], np.float64)
        vit["extra%d" % k] = np.array(sorted(extra.items()), np.int64).reshape(-1, 2)
    vit["n"] = np.array([len(viterbi_cases())], np.int64)
    np.savez_compressed(os.path.join(GOLDEN, "ref_alignment2path.npz"), **vit)
    att = {}
    for k, (seed, rej) in enumerate([(71, False), (72, True), (73, False)]):
        c = attention_case(seed, after_rejection=rej)
        ctx, a = run_reference_attention(c)
        att.update({"%s%d" % (key, k): v for key, v in c.items()})
        att["ctx%d" % k], att["a%d" % k] = ctx, a
    att["n"] = np.array([3], np.int64)
    np.savez_compressed(os.path.join(GOLDEN, "ref_attention_call.npz"), **att)
    print("fixtures written to", GOLDEN)


if __name__ == "__main__":
    main()

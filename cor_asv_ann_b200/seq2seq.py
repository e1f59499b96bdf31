"""Drop-in for ``ocrd_cor_asv_ann.lib.seq2seq.Sequence2Sequence`` on the B200 engine.

Mirrors the reference class (/root/reference/ocrd_cor_asv_ann/lib/seq2seq.py:108-179 attributes,
:190 configure, :782 correct_lines, :1020 vectorize_lines, :1121-1174 save/load_config/load_weights,
:1215 decode_batch_greedy, :1288 decode_sequence_greedy, :1356 decode_sequence_beam) for the
inference path: same attribute names, method signatures, return types and error behaviour, so that
``cor-asv-ann-proc`` / ``-eval`` and ``ocrd-cor-asv-ann-process`` can import this class instead.
All arithmetic runs in libcor_b200.so (CUDA, sm_100a) through ``engine.Engine``; there is no CPU
fallback — decoding without a CUDA device raises.  ``evaluate`` runs its string metrics
(``metrics.py``) on host processes beside the GPU.
"""
import logging
import math
import os

import numpy as np

from . import h5lite
from .vectorize import lines_to_sparse, dense_to_sparse
from .pipeline import StringTable, Prefetcher, chunked, greedy_line_score, merge_batches, LineFiles

GAP = '\a'  # reserved for gap in input (seq2seq.py:29)


class _Layer:
    """What ``scripts/train.py:88`` and weight I/O touch of a Keras layer."""

    def __init__(self, owner, name):
        self._owner, self.name = owner, name

    def get_weights(self):
        return [np.array(a) for a in self._owner._weights[self.name]]

    def set_weights(self, weights):
        self._owner._weights[self.name] = [np.asarray(a, np.float32) for a in weights]
        self._owner._engine = None


class _Model:
    """Stand-in for the Keras ``Model`` objects callers hold on to (``.layers``, ``.name``)."""

    def __init__(self, owner, name, layer_names, predict=None):
        self.name = name
        self.layers = [_Layer(owner, n) for n in layer_names]
        self._predict = predict

    def get_layer(self, name):
        for layer in self.layers:
            if layer.name == name:
                return layer
        raise ValueError('No such layer: ' + name)

    def predict_on_batch(self, x):
        if self._predict is None:
            raise NotImplementedError('%s.predict_on_batch: use the Sequence2Sequence decoders' % self.name)
        return self._predict(x)


class Node:
    """Result view of one beam hypothesis (the trie itself lives on the GPU, csrc/beam.cu)."""

    def __init__(self, value, prob, cum_cost, length, alignment):
        self.value, self.prob, self.cum_cost, self.length, self.alignment = value, prob, cum_cost, length, alignment


class Sequence2Sequence(object):
    '''Sequence to sequence (character-level) error correction, B200 engine behind the reference API.'''

    def __init__(self, logger=None, progbars=True):
        ### model parameters (seq2seq.py:108-132)
        self.batch_size = 256
        self.stateful = False
        self.width = 512
        self.depth = 2
        self.mapping = ({'': 0}, {0: ''})
        self.voc_size = 1
        self.residual_connections = False
        self.deep_bidirectional_encoder = False
        self.bridge_dense = False
        ### training parameters (:134-158) — kept for callers that set them
        self.epochs = 100
        self.lm_loss = False
        self.lm_predict = False
        self.scheduled_sampling = None
        self.dropout = 0.2
        ### beam decoder inference parameters (:160-169)
        self.rejection_threshold = 0.3
        self.beam_width_in = 15
        self.beam_threshold_in = 0.2
        self.beam_width_out = 16
        ### runtime variables (:171-179)
        self.logger = logger or logging.getLogger(__name__)
        self.graph = None
        self.encoder_decoder_model = None
        self.encoder_model = None
        self.decoder_model = None
        from .metrics import Alignment
        self.aligner = Alignment(0, logger=self.logger)  # aligner (for training) with internal state (:177)
        self.progbars = progbars
        self.status = 0  # empty / configured / trained?
        ### engine
        self.device = 0
        self.gemm_mode = 0
        self.beam_budget_bytes = 8 << 30
        self.superbatch = 64     # reference batches per engine call in predict() (16 384 lines at batch_size 256)
        self.prefetch = True     # read + encode the next super-batch in a worker thread
        self.metric_workers = min(8, os.cpu_count() or 1)   # host processes for evaluate()'s string alignments
        self._weights = {}
        self._engine = None

    def __repr__(self):
        return (__name__ +
                " (width: %d)" % self.width +
                " (depth: %d)" % self.depth +
                " (chars: %d)" % self.voc_size +
                " (attention)" +
                (" (stateful)" if self.stateful else " (stateless)") +
                " status: %s" % ("empty" if self.status < 1 else "configured" if self.status < 2 else "trained"))

    # ------------------------------------------------------------------ topology
    @property
    def _enc_dim(self):
        return 2 * self.width if (self.depth == 1 or self.deep_bidirectional_encoder) else self.width

    def _layer_shapes(self):
        """Keras layer name -> list of weight shapes, in Keras weight order."""
        from .synthetic import layer_shapes
        return layer_shapes(self.width, self.depth, self.voc_size, self.deep_bidirectional_encoder, self.bridge_dense)

    def _init_weights(self, rng, variant="K"):
        """Keras default initialisers (seq2seq.py:240: embedding N(0, 0.001); glorot-uniform kernels;
        orthogonal recurrent kernels; zero biases with unit forget gate)."""
        from .synthetic import random_weights
        return random_weights(self._layer_shapes(), self.width, self.depth, rng, variant)

    def configure(self, batch_size=None):
        '''Define encoder and decoder models for the configured parameters (seq2seq.py:190-489).

        The Keras graphs become a weight store plus the CUDA engine; weights start from the
        reference's initialisers and are normally replaced by ``load_weights``.'''
        if batch_size:
            self.batch_size = batch_size
        self.logger.info('using B200 CUDA engine to compile %s model of depth %d width %d size %d with attention',
                         'stateful' if self.stateful else 'stateless', self.depth, self.width, self.voc_size)
        self._weights = self._init_weights(np.random.default_rng())
        self._engine = None
        names = list(self._layer_shapes())
        enc = [n for n in names if n.startswith(('char_input', 'encoder_lstm', 'bridge'))]
        self.encoder_decoder_model = _Model(self, 'encoder_decoder_model', names)
        self.encoder_model = _Model(self, 'encoder_model', enc, predict=self._predict_encoder)
        self.decoder_model = _Model(self, 'decoder_model', [n for n in names if n not in enc or n.startswith('char_input')],
                                    predict=self._predict_decoder)
        self.graph = object()
        self.status = 1

    def _resync_decoder(self):
        """The reference copies the attention layer into a second instance (:526-528); one weight store
        makes this a no-op."""

    # ------------------------------------------------------------------ engine
    def _get_engine(self):
        if self._engine is None:
            from .engine import Engine
            topo = dict(width=int(self.width), depth=int(self.depth), voc_size=int(self.voc_size),
                        residual_connections=bool(self.residual_connections),
                        deep_bidirectional_encoder=bool(self.deep_bidirectional_encoder),
                        bridge_dense=bool(self.bridge_dense), lm_predict=bool(self.lm_predict), window_width=5)
            eos = self.mapping[0].get('\n', 1)
            self._engine = Engine(topo, self._weights, device=self.device, eos_index=eos, gemm_mode=self.gemm_mode)
            self._engine_lm = bool(self.lm_predict)
        elif self._engine_lm != bool(self.lm_predict):
            self._engine = None
            return self._get_engine()
        return self._engine

    # ------------------------------------------------------------------ model files
    @staticmethod
    def _weight_names(layer, count, depth):
        """Dataset names the way Keras names the variables (loading is positional, names are cosmetic)."""
        three = ['kernel:0', 'recurrent_kernel:0', 'bias:0']
        if layer == 'char_input_projection':
            return ['char_input_projection/kernel:0']
        if layer == 'attention_dense':
            return ['attention_dense/kernel:0']
        if layer.startswith('bridge'):
            return [layer + '/kernel:0', layer + '/bias:0']
        if layer == 'decoder_lstm_%d' % depth:
            return [layer + '/' + n for n in ['W_a:0', 'v_a:0', 'b_UW:0', 'b_v:0'] + three]
        if count == 6:
            return [layer + '/forward_' + layer + '/' + n for n in three] + \
                   [layer + '/backward_' + layer + '/' + n for n in three]
        return [layer + '/' + n for n in three]

    def save(self, filename):
        '''Save model weights and configuration parameters (seq2seq.py:1121-1141): Keras
        ``save_weights`` layout plus the ``/config`` group.'''
        assert self.status > 1  # already trained
        self.logger.info('Saving model under "%s"', filename)
        self._save_weights(filename, with_config=True)

    def _save_weights(self, filename, with_config=False):
        """Keras ``save_weights`` layout (what ModelCheckpoint(save_weights_only=True) writes per epoch, seq2seq.py:618),
        plus the reference's ``/config`` group for ``save``."""
        root = h5lite.Group()
        names = list(self._layer_shapes())
        root.attrs['layer_names'] = np.array([n.encode('utf8') for n in names])
        root.attrs['backend'] = np.bytes_(b'tensorflow')
        root.attrs['keras_version'] = np.bytes_(b'2.3.1')
        for layer in names:
            group = root.require_group(layer)
            wnames = self._weight_names(layer, len(self._weights[layer]), self.depth)
            group.attrs['weight_names'] = np.array([n.encode('utf8') for n in wnames])
            for wname, arr in zip(wnames, self._weights[layer]):
                parts = wname.split('/')
                sub = group.require_group('/'.join(parts[:-1])) if len(parts) > 1 else group
                sub[parts[-1]] = np.asarray(arr, np.float32)
        if not with_config:
            h5lite.write(filename, root)
            return
        config = root.require_group('config')
        config['width'] = np.array(int(self.width))
        config['depth'] = np.array(int(self.depth))
        config['stateful'] = np.array(bool(self.stateful))
        config['residual_connections'] = np.array(bool(self.residual_connections))
        config['deep_bidirectional_encoder'] = np.array(bool(self.deep_bidirectional_encoder))
        config['bridge_dense'] = np.array(bool(self.bridge_dense))
        config['mapping'] = np.fromiter((ord(self.mapping[1][i]) if i in self.mapping[1] and self.mapping[1][i] else 0
                                         for i in range(self.voc_size)), dtype=np.uint32)
        h5lite.write(filename, root)

    def load_config(self, filename):
        '''Load parameters to prepare configuration/compilation (seq2seq.py:1143-1162).'''
        config = h5lite.read(filename)['config']
        self.width = int(config['width'][()])
        self.depth = int(config['depth'][()])
        self.stateful = bool(config['stateful'][()])
        self.residual_connections = bool(config['residual_connections'][()]) \
            if 'residual_connections' in config else False  # old default
        self.deep_bidirectional_encoder = bool(config['deep_bidirectional_encoder'][()]) \
            if 'deep_bidirectional_encoder' in config else False  # old default
        self.bridge_dense = bool(config['bridge_dense'][()]) if 'bridge_dense' in config else False  # old default
        c_i = dict((chr(c), i) if c > 0 else ('', 0) for i, c in enumerate(config['mapping'][()]))
        i_c = dict((i, chr(c)) if c > 0 else (0, '') for i, c in enumerate(config['mapping'][()]))
        self.mapping = (c_i, i_c)
        self.voc_size = len(c_i)

    @staticmethod
    def _from_cudnn(weights, W):
        """Keras ``_convert_rnn_weights`` CuDNNLSTM -> LSTM: a layer saved from a GPU-trained model has
        bias length 8W and per-gate transposed kernels."""
        out = []
        for i in range(0, len(weights), 3):
            k, r, b = weights[i:i + 3]
            if b.shape != (8 * W,):
                out += [k, r, b]
                continue
            k = np.hstack([g.T.reshape(g.shape, order='F') for g in np.hsplit(k, 4)])
            r = np.hstack([g.T for g in np.hsplit(r, 4)])
            out += [k, r, b.reshape((2, -1)).sum(axis=0)]
        return out

    def _read_layer_weights(self, root, by_name_only=True):
        if 'layer_names' not in root.attrs and 'model_weights' in root:
            root = root['model_weights']
        found = {}
        for raw in root.attrs.get('layer_names', []):
            layer = raw.decode('utf8') if isinstance(raw, bytes) else str(raw)
            if layer not in root:
                continue
            group = root[layer]
            wnames = [n.decode('utf8') if isinstance(n, bytes) else str(n) for n in group.attrs.get('weight_names', [])]
            if wnames:
                found[layer] = [np.asarray(group.get_path(n), np.float32) for n in wnames]
        return found

    def load_weights(self, filename):
        '''Load weights into the configured/compiled model (seq2seq.py:1164-1174): by layer name,
        positional within the layer; layers of other shape raise like Keras does.'''
        assert self.status > 0  # already compiled
        self.logger.info('Loading model from "%s"', filename)
        found = self._read_layer_weights(h5lite.read(filename))
        W = int(self.width)
        for layer, shapes in self._layer_shapes().items():
            if layer not in found:
                continue
            ws = found[layer]
            if 'lstm' in layer:
                off = 4 if layer == 'decoder_lstm_%d' % self.depth else 0
                ws = ws[:off] + self._from_cudnn(ws[off:], W)
            if len(ws) != len(shapes):
                raise ValueError('Layer %s expects %d weight(s), but the saved weights have %d element(s)'
                                 % (layer, len(shapes), len(ws)))
            for a, shape in zip(ws, shapes):
                if tuple(a.shape) != tuple(shape):
                    raise ValueError('Layer %s: weight shape %s incompatible with saved shape %s'
                                     % (layer, shape, a.shape))
            self._weights[layer] = [np.ascontiguousarray(a, np.float32) for a in ws]
        self._engine = None
        self._resync_decoder()
        self.status = 2

    def load_transfer_weights(self, filename):
        '''Load weights from another model into the matching layers (seq2seq.py:1176-1213):
        layers whose shapes do not match are skipped (``skip_mismatch=True``).'''
        assert self.status > 0  # already compiled
        assert self.depth > 1
        root = h5lite.read(filename)
        if 'config' in root:
            config = root['config']
            c_i = dict((chr(c), i) if c > 0 else ('', 0) for i, c in enumerate(config['mapping'][()]))
            i_c = dict((i, chr(c)) if c > 0 else (0, '') for i, c in enumerate(config['mapping'][()]))
            if len(c_i) != self.voc_size:
                self.mapping, self.voc_size = (c_i, i_c), len(c_i)
                self._weights.update({k: v for k, v in self._init_weights(np.random.default_rng()).items()
                                      if k == 'char_input_projection'})
            else:
                self.mapping = (c_i, i_c)
        self.logger.info('Transferring model from "%s"', filename)
        found = self._read_layer_weights(root)
        W = int(self.width)
        for layer, shapes in self._layer_shapes().items():
            if layer not in found:
                continue
            ws = found[layer]
            off = 0
            if layer == 'decoder_lstm_%d' % self.depth:
                # the LM donor has no attention wrapper in its top layer: only the cell weights transfer
                off = 4
                ws = list(self._weights[layer][:4]) + (ws[4:] if len(ws) == 7 else ws)
            if 'lstm' in layer:
                ws = ws[:off] + self._from_cudnn(ws[off:], W)
            if len(ws) != len(shapes) or any(tuple(a.shape) != tuple(s) for a, s in zip(ws, shapes)):
                self.logger.warning('skipping layer %s due to mismatch', layer)
                continue
            self._weights[layer] = [np.ascontiguousarray(a, np.float32) for a in ws]
        self._engine = None
        self.status = 1

    def map_files(self, filenames):
        """Character inventory and line count of the files (seq2seq.py:554-590); extends the mapping."""
        from .training import map_files
        return map_files(self, filenames)

    def _reconfigure_for_mapping(self):
        """seq2seq.py:499-525: the embedding grows with the mapping; weights of the characters known before keep their
        rows, every other layer keeps its weights (incremental training), or everything is initialised afresh."""
        old, had = self._weights, self.status >= 2
        old_rows = old['char_input_projection'][0].shape[0] if old else 0
        if old_rows >= self.voc_size:
            return
        self.configure()
        if had:
            self.logger.warning('transferring weights from previous model with only %d character types', old_rows)
            for layer, arrs in old.items():
                if layer == 'char_input_projection':
                    self._weights[layer][0][:old_rows] = arrs[0]
                else:
                    self._weights[layer] = arrs

    def train(self, filenames, val_filenames=None):
        '''train model on given text files (seq2seq.py:592-650): epochs of data-parallel steps on the CUDA trainer
        (csrc/train.cu) with early stopping on the validation loss, see ``training.train``.'''
        from .training import train
        return train(self, filenames, val_filenames)

    def evaluate(self, filenames, fast=False, normalization='historic_latin', charmap=None, gt_level=1,
                 confusion=10, histogram=True):
        '''evaluate model on text files (seq2seq.py:651-754): character and word error rates of the OCR input
        and of the greedy and beamed predictions against the ground truth, confusion tables, histograms and
        perplexities, logged like the reference does.

        Same batches, lines and numbers as the reference's loop.  `superbatch` reference batches share one
        engine call, and the string metrics of a group (difflib alignments, pure Python) run on
        `metric_workers` host processes while the GPU decodes the next group; they are folded into the running
        statistics in line order, so the order-dependent recurrences give the reference's values.  The final
        figures are also kept in ``self.last_evaluation``.'''
        assert self.status == 2
        from .metrics import Alignment, Edits, LineScorer
        kinds = ('origin', 'greedy', 'beamed')
        counts = {lv + '_' + k: Edits(self.logger, histogram=histogram and lv == 'c') for lv in 'cw' for k in kinds}
        aligners = {lv + '_' + k: Alignment(0, logger=self.logger, confusion=confusion > 0 and lv == 'c')
                    for lv in 'cw' for k in kinds}
        scorer = LineScorer(self.metric_workers)

        def fold(group, greedy, beamed, results):
            rows = iter(results())
            for batch, res_g, res_b in zip(group, greedy, beamed):
                lines_source, _, lines_target, _ = batch
                for j in range(len(lines_source)):
                    if not lines_source[j] or not lines_target[j]:
                        continue  # from partially filled batch
                    self.logger.info('Source input              : %s', lines_source[j].rstrip(u'\n'))
                    self.logger.info('Target output             : %s', lines_target[j].rstrip(u'\n'))
                    self.logger.info('Target prediction (greedy): %s [%.2f]', res_g[0][j].rstrip(u'\n'), res_g[2][j])
                    self.logger.info('Target prediction (beamed): %s [%.2f]', res_b[0][j].rstrip(u'\n'), res_b[2][j])
                    row = next(rows)
                    for name in ('c_origin', 'c_greedy', 'c_beamed', 'w_origin', 'w_greedy', 'w_beamed'):
                        dist, length, seq1, seq2, conf = row[name]
                        if conf:
                            aligners[name].count_confusion(conf)
                        counts[name].add(dist, length, seq1, seq2)
                counts['c_greedy'].score += sum(res_g[2])
                counts['c_beamed'].score += sum(res_b[2])

        batches = self.gen_lines(filenames, False, charmap=charmap)
        groups = chunked(batches, max(int(self.superbatch), 1))
        if self.prefetch:
            groups = Prefetcher(groups, depth=2)
        waiting = None
        try:
            for group in groups:
                pairs = [(b[0], b[1]) for b in group]
                greedy = self.correct_batches(pairs, fast=fast, greedy=True, want_probs=False)
                beamed = greedy if fast else self.correct_batches(pairs, fast=False, greedy=False, want_probs=False)
                if waiting is not None:
                    fold(*waiting)          # the previous group's alignments ran while this one was decoded
                jobs = [(b[0][j], b[2][j], g[0][j], m[0][j], normalization, gt_level, confusion > 0)
                        for b, g, m in zip(group, greedy, beamed) for j in range(len(b[0])) if b[0][j] and b[2][j]]
                waiting = (group, greedy, beamed, scorer.submit(jobs))
            if waiting is not None:
                fold(*waiting)
        finally:
            scorer.close()

        c_o, c_g, c_b = counts['c_origin'], counts['c_greedy'], counts['c_beamed']
        w_o, w_g, w_b = counts['w_origin'], counts['w_greedy'], counts['w_beamed']
        self.logger.info('finished %d lines', c_o.length)
        if confusion > 0:
            self.logger.info('OCR    confusion: %s', aligners['c_origin'].get_confusion(confusion))
            self.logger.info('greedy confusion: %s', aligners['c_greedy'].get_confusion(confusion))
            self.logger.info('beamed confusion: %s', aligners['c_beamed'].get_confusion(confusion))
        if histogram:
            self.logger.info('OCR    histogram: %s', repr(c_o.hist()))
            self.logger.info('greedy histogram: %s', repr(c_g.hist()))
            self.logger.info('beamed histogram: %s', repr(c_b.hist()))
        self.logger.info('ppl greedy: %.3f', math.exp(c_g.score / c_g.length))
        self.logger.info('ppl beamed: %.3f', math.exp(c_b.score / c_b.length))
        self.logger.info("CER OCR:    %.3f±%.3f", c_o.mean, math.sqrt(c_o.varia))
        self.logger.info("CER greedy: %.3f±%.3f", c_g.mean, math.sqrt(c_g.varia))
        self.logger.info("CER beamed: %.3f±%.3f", c_b.mean, math.sqrt(c_b.varia))
        self.logger.info("WER OCR:    %.3f±%.3f", w_o.mean, math.sqrt(w_o.varia))
        self.logger.info("WER greedy: %.3f±%.3f", w_g.mean, math.sqrt(w_g.varia))
        self.logger.info("WER beamed: %.3f±%.3f", w_b.mean, math.sqrt(w_b.varia))
        self.last_evaluation = {name: dict(mean=c.mean, varia=c.varia, length=c.length, steps=c.steps, score=c.score)
                                for name, c in counts.items()}
        self.last_evaluation['confusion'] = {k: aligners['c_' + k].get_confusion(confusion) for k in kinds} if confusion > 0 else {}

    # ------------------------------------------------------------------ data
    def gen_lines(self, filenames, repeat=True, split=None, train=False, unsupervised=False, charmap=None):
        """Generate batches of lines from the given files (seq2seq.py:919-1018): 4-tuples (sources, source
        confidences or None, targets, filenames) of ``batch_size`` lines in file order.  With ``repeat`` the
        files are read again and again, ``False`` marks the end of every pass and the incomplete last batch of
        a pass is dropped; without, there is one pass whose last batch is padded with empty lines.
        Reading, normalisation and batching live in ``pipeline.LineFiles``."""
        def too_bad(source, target):                    # training skips pairs the OCR got too wrong (:979-987)
            self.aligner.set_seqs(source, target)
            return self.aligner.is_bad()
        while True:
            reader = LineFiles(filenames, split=split, train=train, unsupervised=unsupervised, charmap=charmap,
                               reject=too_bad if train and self.aligner is not None else None)
            yield from reader.batches(self.batch_size, pad_last=not repeat)
            if not repeat:
                return
            yield False

    def predict(self, filenames, fast=False, greedy=False, charmap=None):
        '''apply model on text files (seq2seq.py:756-780): yields (filenames, lines, scores) per batch.

        Same batches, same order and same results as the reference's loop, but `superbatch` reference
        batches share one engine call (correct_batches) and the next super-batch is read and encoded in
        a worker thread while the GPU decodes the current one.'''
        assert self.status == 2
        batches = self.gen_lines(filenames, repeat=False, unsupervised=True, charmap=charmap)
        groups = chunked(batches, max(int(self.superbatch), 1))
        if fast and self.prefetch:
            yield from self._predict_pipelined(groups)
            return
        if self.prefetch:
            groups = Prefetcher(groups, depth=2)
        for group in groups:
            results = self.correct_batches([(b[0], b[1]) for b in group], fast=fast, greedy=greedy, want_probs=False)
            for batch, (lines_result, _, scores_result, _) in zip(group, results):
                yield (batch[3], lines_result, scores_result)

    def _predict_pipelined(self, groups):
        """Fast mode as three overlapping stages: (1) read + index-encode + merge the next super-batch, (2) the engine
        call (the only stage that touches the GPU; the ctypes call releases the GIL), (3) strings and per-batch
        results of the previous one -- each in its own thread, two super-batches of slack between neighbours.
        Same batches, same order, same results as correct_batches per group."""
        def encode():
            for group in groups:
                todo = [(b, batch[0], batch[1]) for b, batch in enumerate(group) if batch[0]]
                if not todo:
                    yield group, todo, None
                    continue
                enc = [lines_to_sparse(self.mapping[0], lines, conf, self.logger) for _, lines, conf in todo]
                yield group, todo, merge_batches(enc)

        def decode(items):
            eng = self._get_engine()
            pool, turn = [None] * 6, 0          # result buffers in flight: 2 queued + 1 being read + 1 being written, and slack
            for group, todo, merged in items:
                if not todo:
                    yield group, todo, None, None
                    continue
                idx, val, row_T, offsets = merged
                want = (idx.shape[0], 2 * idx.shape[1])
                out = pool[turn] if pool[turn] is not None and tuple(pool[turn][0].shape) == want else None
                out = eng.correct_greedy_host(idx, val, row_T, want_align=False, out=out)
                pool[turn] = out
                turn = (turn + 1) % len(pool)
                yield group, todo, (out[0].numpy(), out[2].numpy(), out[3].numpy()), offsets

        for group, todo, arrays, offsets in Prefetcher(decode(Prefetcher(encode(), depth=2)), depth=2):
            results = [([], []) for _ in group]
            if todo:
                chars, scores, lens = arrays
                lines = self._strings(chars, lens)
                out_scores = np.where(lens > 0, scores, 0.).tolist()
                for n, (b, _, _) in enumerate(todo):
                    lo, hi = offsets[n], offsets[n + 1]
                    results[b] = (lines[lo:hi], out_scores[lo:hi])
            for batch, (lines_result, scores_result) in zip(group, results):
                yield (batch[3], lines_result, scores_result)

    def vectorize_lines(self, encoder_input_sequences, decoder_input_sequences, encoder_conf_sequences=None):
        '''Convert a batch of source and target sequences to dense arrays (seq2seq.py:1020-1119).'''
        assert len(encoder_input_sequences) == len(decoder_input_sequences)
        batch_size = len(encoder_input_sequences)
        idx, val = lines_to_sparse(self.mapping[0], encoder_input_sequences, encoder_conf_sequences, self.logger)
        B, T, k = idx.shape
        enc = np.zeros((batch_size, T, self.voc_size), dtype=np.float32 if encoder_conf_sequences else np.uint32)
        for a in range(k):
            b, t = np.nonzero(idx[:, :, a] >= 0)
            enc[b, t, idx[b, t, a]] = 1 if val is None else val[b, t, a]
        Td = max(map(len, decoder_input_sequences))
        dec_in = np.zeros((batch_size, Td + 1, self.voc_size), dtype=np.uint32)
        dec_out = np.zeros((batch_size, Td + 1, self.voc_size), dtype=np.uint32)
        for i, dec_seq in enumerate(decoder_input_sequences):
            for j, char in enumerate(dec_seq):
                if char not in self.mapping[0]:
                    if char != GAP:
                        self.logger.error('unmapped character "%s" at decoder input sequence %d', char, i)
                    v = 0
                else:
                    v = self.mapping[0][char]
                dec_in[i, j + 1, v] = 1
                dec_out[i, j, v] = 1
        weights = np.ones(dec_out.shape[:-1], dtype=np.float32)
        weights[np.all(dec_out == 0, axis=2)] = 0.
        if self.lm_loss:
            dec_out, weights = [dec_out, dec_out], [weights, weights]
        return enc, dec_in, dec_out, weights

    # ------------------------------------------------------------------ decoding
    def _row_T(self, idx):
        return np.full(idx.shape[0], idx.shape[1], np.int32)    # one call = one reference batch: T is its padded length

    def _predict_encoder(self, encoder_input_data):
        """encoder_model.predict_on_batch (seq2seq.py:403-406): [enc_out (B,T,D)] + [h_n, c_n]*depth + [a0 (B,T)]."""
        import torch
        eng = self._get_engine()
        idx, val = dense_to_sparse(encoder_input_data)
        dev = eng.device
        enc_out, states = eng.encode(torch.as_tensor(idx, device=dev),
                                     None if val is None else torch.as_tensor(val, device=dev),
                                     torch.as_tensor(self._row_T(idx), device=dev))
        out = [enc_out.permute(1, 0, 2).contiguous().cpu().numpy()]
        out += [s.cpu().numpy() for s in states]
        out.append(np.zeros(idx.shape[:2], np.float32))
        return out

    def _preset_from(self, encoder_outputs):
        """[enc_out (B,T,D)] + [h_n, c_n]*depth + [a0] (the encoder model's output list, or [j:j+1] slices of it)
        -> the engine's preset layouts: enc_out [T,B,D], states [2d,B,W]."""
        import torch
        enc_out = np.ascontiguousarray(np.asarray(encoder_outputs[0], np.float32).transpose(1, 0, 2))
        states = np.stack([np.asarray(s, np.float32) for s in encoder_outputs[1:1 + 2 * int(self.depth)]])
        return torch.from_numpy(enc_out), torch.from_numpy(states)

    def _predict_decoder(self, inputs):
        """decoder_model.predict_on_batch (seq2seq.py:477-480): [p (B,1,V), enc_out (B or 1,T,D)] + [h_1, c_1, ..., h_d,
        c_d, a] -> [scores (B,1,V)] (+ [lm scores]) + the new states, one cor_decode_step on the engine."""
        import torch
        eng = self._get_engine()
        dev = eng.device
        p = np.asarray(inputs[0], np.float32)
        attended = np.asarray(inputs[1], np.float32)
        states = [np.asarray(s, np.float32) for s in inputs[2:]]
        R, d = p.shape[0], int(self.depth)
        L, T = attended.shape[0], attended.shape[1]
        assert L in (1, R), "attended sequence must have batch 1 or the batch of the states"
        # the attended lines enter as a preset encoding (their encoder states are irrelevant for a single step)
        if L == 1 and R > 1:        # the beam's broadcast form (:1428-1429): every hypothesis row gets the one line
            attended, L = np.repeat(attended, R, axis=0), R
        eng.set_encoding(torch.from_numpy(np.ascontiguousarray(attended.transpose(1, 0, 2))),
                         torch.zeros((2 * d, L, int(self.width))))
        try:
            idx = torch.ones((L, T, 1), dtype=torch.int32, device=dev)
            eng.encode(idx, None, torch.full((L,), T, dtype=torch.int32, device=dev), want_outputs=False)
        finally:
            eng.set_encoding()
        row_line = None
        h_in = torch.as_tensor(np.stack(states[0:2 * d:2]), device=dev)
        c_in = torch.as_tensor(np.stack(states[1:2 * d:2]), device=dev)
        p_out, lm_out, h_out, c_out, a_out = eng.decode_step(torch.as_tensor(p[:, -1, :], device=dev), h_in, c_in,
                                                             torch.as_tensor(states[2 * d], device=dev), row_line)
        out = [p_out.cpu().numpy()[:, None, :]]
        if self.lm_predict:
            out.append(lm_out.cpu().numpy()[:, None, :])
        for n in range(d):
            out += [h_out[n].cpu().numpy(), c_out[n].cpu().numpy()]
        out.append(a_out.cpu().numpy())
        return out

    def _strings(self, chars, lens):
        table = getattr(self, '_string_table', None)
        if table is None or table.i2c is not self.mapping[1]:
            table = self._string_table = StringTable(self.mapping[1], self.voc_size)
        return table.strings(chars, lens)

    def _greedy_sparse(self, idx, val, want_align=True, row_T=None, want_probs=True):
        eng = self._get_engine()
        if row_T is None:
            row_T = self._row_T(idx)
        chars, probs, scores, lens, align = eng.correct_greedy_host(idx, val, row_T, want_align=want_align)
        chars, probs, scores, lens = chars.numpy(), probs.numpy(), scores.numpy(), lens.numpy()
        B = idx.shape[0]
        lines = self._strings(chars, lens)
        # predict() drops the probabilities: hand out array rows then instead of 2T Python floats per line
        out_probs = [list(probs[j, :lens[j]]) for j in range(B)] if want_probs else [probs[j, :lens[j]] for j in range(B)]
        out_scores = np.where(lens > 0, scores, 0.).tolist()
        if align is not None:
            align = align.numpy()
            # one copy per line, then S row views of it (the reference returns a list of S arrays (T,))
            alignments = [list(np.ascontiguousarray(align[:lens[j], j, :row_T[j]])) for j in range(B)]
        else:
            alignments = [[] for _ in range(B)]
        return lines, out_probs, out_scores, alignments

    def correct_lines(self, lines, conf=None, fast=True, greedy=True):
        '''apply correction model on text strings (seq2seq.py:782-842).

        Return a 4-tuple of the corrected lines, probability lists, perplexity scores, and
        input-output alignments.'''
        assert not fast or greedy, "cannot decode in fast mode with beam search enabled"
        if not lines:
            return [], [], [], []
        idx, val = lines_to_sparse(self.mapping[0], lines, conf, self.logger)
        return self._correct_sparse(lines, idx, val, None, fast, greedy, True)

    def _correct_sparse(self, lines, idx, val, row_T, fast, greedy, want_align, want_probs=True):
        if fast:
            return self._greedy_sparse(idx, val, want_align, row_T, want_probs)
        if greedy:
            # decode_sequence_greedy per line (:817-819) picks the same characters as the batch decoder —
            # "argmax, and if '' wins the argmax of the rest" == argmax over indices >= 1 — from the same
            # batch encoding; only the score has its own accumulation there (float32 logarithms, :1330, :1341)
            out_lines, out_probs, _, alignments = self._greedy_sparse(idx, val, want_align, row_T, want_probs)
            out_scores = []
            for j, line in enumerate(lines):
                if not line:
                    out_lines[j], out_probs[j], alignments[j] = '', [], []
                    out_scores.append(0)
                    continue
                out_scores.append(greedy_line_score(out_probs[j]) if out_lines[j] else 0)
                out_lines[j] = out_lines[j].replace(GAP, '')
            return out_lines, out_probs, out_scores, alignments
        return self._beam_sparse(lines, idx, val, want_align, row_T)

    def correct_batches(self, batches, fast=True, greedy=True, want_align=False, want_probs=True):
        '''correct_lines over several reference batches in ONE engine call.

        `batches`: list of (lines, conf) pairs, each what one correct_lines call would get (e.g. the
        256-line batches of gen_lines, or the pages of a document).  Every line is decoded with the
        padded length of ITS batch (the reference does not mask padding, so T is part of the input),
        hence the result of batch b equals correct_lines(*batches[b]) — without the soft alignments
        unless `want_align`; with `want_probs=False` the probability entries are float32 array rows
        rather than lists.  Return a list of 4-tuples, one per batch.'''
        assert not fast or greedy, "cannot decode in fast mode with beam search enabled"
        todo = [(b, lines, conf) for b, (lines, conf) in enumerate(batches) if lines]
        results = [([], [], [], []) for _ in batches]
        if not todo:
            return results
        encoded = [lines_to_sparse(self.mapping[0], lines, conf, self.logger) for _, lines, conf in todo]
        idx, val, row_T, offsets = merge_batches(encoded)
        all_lines = [line for _, lines, _ in todo for line in lines]
        merged = self._correct_sparse(all_lines, idx, val, row_T, fast, greedy, want_align, want_probs)
        for n, (b, _, _) in enumerate(todo):
            lo, hi = offsets[n], offsets[n + 1]
            results[b] = tuple(part[lo:hi] for part in merged)
        return results

    def correct_lines_paths(self, lines, conf=None, fast=True, greedy=True, input_lens=None):
        '''correct_lines plus the hard re-alignment path of every line, made on the device.

        Replaces the pair "correct_lines, then per line _alignment2path(alignment, input_line_len,
        len(output_line), 1. / voc_size)" of the OCR-D wrapper (wrapper/transcode.py:111-127,
        279-349): the soft alignments never leave the GPU; only T+1 ints and a float per line do.
        `input_lens`: input_line_len per line (transcode.py:123; default: number of input positions).
        Return a 5-tuple: lines, probability lists, scores, realignment dicts {input pos: output pos},
        distances — the first three exactly as correct_lines returns them.'''
        import torch
        assert not fast or greedy, "cannot decode in fast mode with beam search enabled"
        if not lines:
            return [], [], [], [], []
        eng = self._get_engine()
        dev = eng.device
        idx, val = lines_to_sparse(self.mapping[0], lines, conf, self.logger)
        B, T, _ = idx.shape
        d_idx = torch.as_tensor(idx, device=dev)
        d_val = None if val is None else torch.as_tensor(val, device=dev)
        d_T = torch.as_tensor(self._row_T(idx), device=dev)
        i2c = self.mapping[1]
        if input_lens is None:
            input_lens = [len(line) for line in lines]
        if greedy:
            eng.encode(d_idx, d_val, d_T, want_align=True)
            chars, probs, scores, lens, align = eng.greedy()
            layout = "greedy"
        else:
            bp = eng.beam_params(batch_size=min(int(self.batch_size), 256), beam_width_in=self.beam_width_in,
                                 beam_threshold_in=self.beam_threshold_in, beam_width_out=self.beam_width_out,
                                 rejection_threshold=self.rejection_threshold)
            out = (torch.empty((B, 2 * T), dtype=torch.int32, device=dev), torch.empty((B, 2 * T), dtype=torch.float32, device=dev),
                   torch.empty((B,), dtype=torch.float64, device=dev), torch.empty((B,), dtype=torch.int32, device=dev),
                   torch.empty((B, 2 * T, T), dtype=torch.float32, device=dev), torch.empty((B, 8), dtype=torch.int64, device=dev))
            chars, probs, scores, lens, align, stats = eng.beam_device(d_idx, d_val, d_T, bp,
                                                                       budget_bytes=self.beam_budget_bytes, out=out)
            self.last_beam_stats = stats.cpu().numpy()
            layout = "beam"
        chars, probs, scores, lens = chars.cpu().numpy(), probs.cpu().numpy(), scores.cpu().numpy(), lens.cpu().numpy()
        out_lines, out_probs, out_scores, out_len = [], [], [], []
        for j, input_line in enumerate(lines):
            n = int(lens[j])
            if not input_line:
                line, prob, score, jl = '', [], 0, 0
            elif n > 0:
                line = ''.join(i2c[int(v)] for v in chars[j, :n])
                prob = list(probs[j, :n])
                if fast or not greedy:
                    score = float(scores[j])
                else:                       # decode_sequence_greedy's own accumulation (:1318, :1341)
                    score = greedy_line_score(prob)
                if not fast:
                    line = line.replace(GAP, '')
                jl = len(line)
            else:                           # beam fallback (:826-836): input returned, identity alignment
                self.logger.error('cannot beam-decode input line %d: "%s"', j, input_line)
                if isinstance(input_line[0], tuple):
                    line = ''.join(chunk[0] for chunk in input_line)
                if isinstance(input_line[0], list):
                    line = ''.join(chunk[0][0] if chunk else '' for chunk in input_line)
                else:
                    line = input_line
                prob, score = [1.0] * len(line), 0
                line = line.replace(GAP, '')
                jl = -len(prob)
            out_lines.append(line)
            out_probs.append(prob)
            out_scores.append(score)
            out_len.append(jl)
        path, dist = eng.alignment_paths(align, layout, np.asarray(input_lens, np.int32),
                                         np.asarray(out_len, np.int32), 1. / self.voc_size)
        path, dist = path.cpu().numpy(), dist.cpu().numpy()
        realignments = []
        for j in range(B):
            i_max = int(input_lens[j])
            if i_max < 1 or out_len[j] == 0:
                realignments.append({})
                continue
            row = path[j, :i_max + 1]
            d = {i_max: int(row[i_max])}                      # same insertion order as the reference's dict
            d.update({int(i): int(row[i]) for i in range(i_max - 1, 0, -1) if row[i] >= 0})
            d[0] = 0
            realignments.append(d)
        return out_lines, out_probs, out_scores, realignments, [float(x) for x in dist]

    def _beam_sparse(self, lines, idx, val, want_align=True, row_T=None):
        eng = self._get_engine()
        if row_T is None:
            row_T = self._row_T(idx)
        bp = eng.beam_params(batch_size=min(int(self.batch_size), 256), beam_width_in=self.beam_width_in,
                             beam_threshold_in=self.beam_threshold_in, beam_width_out=self.beam_width_out,
                             rejection_threshold=self.rejection_threshold)
        chars, probs, scores, lens, align, stats = eng.correct_beam_host(
            idx, val, row_T, bp, want_align=want_align, budget_bytes=self.beam_budget_bytes)
        chars, probs, scores, lens, stats = chars.numpy(), probs.numpy(), scores.numpy(), lens.numpy(), stats.numpy()
        align = None if align is None else align.numpy()
        decoded = self._strings(chars, lens)
        output_lines, output_probs, output_scores, alignments = [], [], [], []
        for j, input_line in enumerate(lines):
            if not input_line:
                line, prob, score, alignment = '', [], 0, []
            elif lens[j] > 0:
                n = int(lens[j])
                line = decoded[j]
                prob = list(probs[j, :n])
                score = float(scores[j])
                alignment = list(np.ascontiguousarray(align[j, :n, :row_T[j]])) if align is not None else []
            else:
                # no final hypothesis: the reference's StopIteration fallback (:826-836)
                self.logger.error('cannot beam-decode input line %d: "%s"', j, input_line)
                if isinstance(input_line[0], tuple):
                    line = ''.join(chunk[0] for chunk in input_line)
                if isinstance(input_line[0], list):
                    line = ''.join(chunk[0][0] if chunk else '' for chunk in input_line)
                else:
                    line = input_line
                prob = [1.0] * len(line)
                score = 0
                alignment = np.eye(len(line)).tolist()
            line = line.replace(GAP, '')
            output_lines.append(line)
            output_probs.append(prob)
            output_scores.append(score)
            alignments.append(alignment)
        self.last_beam_stats = stats
        return output_lines, output_probs, output_scores, alignments

    def decode_batch_greedy(self, encoder_input_data):
        '''Predict from one batch of lines array without alternatives (seq2seq.py:1215-1286).

        Return a 5-tuple of the full output array (for training phase), output strings, output
        probability lists, entropies, and soft alignments.'''
        encoder_input_data = np.asarray(encoder_input_data)
        B, T = encoder_input_data.shape[:2]
        idx, val = dense_to_sparse(encoder_input_data)
        lines, probs, scores, alignments = self._greedy_sparse(idx, val)
        # the reference's first value is the uint32 cast of the fed-back softmax rows, i.e. all zero unless a
        # probability reaches 1.0 (:1237, :1244); only scheduled-sampling training reads it
        decoder_output_data = np.zeros((B, T * 2, self.voc_size), dtype=np.uint32)
        return decoder_output_data, lines, probs, scores, alignments

    def decode_sequence_greedy(self, source_seq=None, encoder_outputs=None):
        '''Predict from one line vector without alternatives (seq2seq.py:1288-1354).

        ``encoder_outputs`` (the encoder model's output list for this line, e.g. ``[e[j:j+1] for e in
        encoder_model.predict_on_batch(batch)]`` as correct_lines does, :817-819) are imported into the engine
        instead of encoding ``source_seq`` again.'''
        if encoder_outputs is None:
            idx, val = dense_to_sparse(np.asarray(source_seq)[None])
            lines, probs, _, alignments = self._greedy_sparse(idx, val)
        else:
            eng = self._get_engine()
            T = np.asarray(encoder_outputs[0]).shape[1]
            if source_seq is not None:
                idx, val = dense_to_sparse(np.asarray(source_seq)[None])
            else:
                idx, val = np.ones((1, T, 1), np.int32), None       # any non-empty line: only the empty-line flag reads it
            eng.set_encoding(*self._preset_from(encoder_outputs))
            try:
                lines, probs, _, alignments = self._greedy_sparse(idx, val)
            finally:
                eng.set_encoding()
        return lines[0], probs[0], greedy_line_score(probs[0]), alignments[0]

    def decode_sequence_beam(self, source_seq=None, encoder_outputs=None):
        '''Predict from one line vector with alternatives (seq2seq.py:1356-1544).

        Generator of (string, probabilities, entropy, alignments) over the final hypotheses of the search, best
        first like the reference's (:1538-1544); the engine hands back the best ``N_BEST`` (64) of them, far more
        than ``beam_width_out`` + 1, after which the search normally stops.  ``encoder_outputs`` (this line's
        slices of an earlier batch encoding, :823-825) are imported instead of encoding ``source_seq`` again.'''
        assert source_seq is not None, "the beam reads its rejection candidates from source_seq (seq2seq.py:1455)"
        idx, val = dense_to_sparse(np.asarray(source_seq)[None])
        eng = self._get_engine()
        bp = eng.beam_params(batch_size=min(int(self.batch_size), 256), beam_width_in=self.beam_width_in,
                             beam_threshold_in=self.beam_threshold_in, beam_width_out=self.beam_width_out,
                             rejection_threshold=self.rejection_threshold, n_best=self.N_BEST)
        if encoder_outputs is not None:
            eng.set_encoding(*self._preset_from(encoder_outputs))
        try:
            chars, probs, scores, lens, align, stats = eng.correct_beam_host(
                idx, val, self._row_T(idx), bp, want_align=True, budget_bytes=self.beam_budget_bytes)
        finally:
            if encoder_outputs is not None:
                eng.set_encoding()
        self.last_beam_stats = stats.numpy()
        i2c = self.mapping[1]
        for r in range(chars.shape[1]):
            n = int(lens[0, r])
            if n <= 0:
                break
            yield (''.join(i2c[int(v)] for v in chars[0, r, :n]), list(probs[0, r, :n].numpy()), float(scores[0, r]),
                   [align[0, r, s].numpy().copy() for s in range(n)])

    N_BEST = 64     # final hypotheses decode_sequence_beam can yield (COR_BEAM_NBEST_MAX)

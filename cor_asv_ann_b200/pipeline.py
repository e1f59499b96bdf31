"""Host side of the line stream around the engine.

The reference's ``predict`` (ocrd_cor_asv_ann/lib/seq2seq.py:756-780) walks ``gen_lines`` one
256-line batch at a time: read, vectorise into a dense (B,T,V) array, decode, yield.  At B200 decode
rates one reference batch is a page-sized call (latency bound, ~13 ms whatever its size), and the
Python loops around it cost more than the GPU work.  Here

* several reference batches share ONE engine call: rows carry the padded length T of the batch they
  came from (``row_T``), so every line sees exactly the zero padding it would see in its own
  256-line batch (the padding is not masked in the reference, it changes results);
* reading + vectorising the next super-batch runs in a worker thread while the GPU decodes the current
  one (the ctypes call releases the GIL);
* result strings are built by one table lookup over the whole (B,S) index array.
"""
import pickle
import queue
import threading
import unicodedata

import numpy as np


def merge_batches(encoded):
    """[(idx [B_b,T_b,k_b], val or None), ...] -> (idx [sum B,T,k], val or None, row_T [sum B], offsets).

    Slots beyond a batch's own T_b / k_b hold index -1 (the engine's "true zero" marker)."""
    B = sum(e[0].shape[0] for e in encoded)
    T = max([e[0].shape[1] for e in encoded] + [1])
    k = max([e[0].shape[2] for e in encoded] + [1])
    soft = any(e[1] is not None for e in encoded)
    idx = np.full((B, T, k), -1, dtype=np.int32)
    val = np.zeros((B, T, k), dtype=np.float32) if soft else None
    row_T = np.empty(B, dtype=np.int32)
    offsets = [0]
    for e_idx, e_val in encoded:
        b, t, kk = e_idx.shape
        o = offsets[-1]
        idx[o:o + b, :t, :kk] = e_idx
        if soft:
            val[o:o + b, :t, :kk] = 1.0 if e_val is None else e_val
            if e_val is None:
                val[o:o + b, :t, :kk][e_idx < 0] = 0.0
        row_T[o:o + b] = t
        offsets.append(o + b)
    return idx, val, row_T, offsets


class StringTable:
    """idx -> char for whole result arrays.  Falls back to a per-character join for alphabets with
    multi-codepoint entries (none in the reference's models, whose alphabet is single characters)."""

    def __init__(self, i2c, voc_size):
        self.i2c = i2c
        self.simple = all(len(i2c.get(i, '')) == 1 for i in range(1, voc_size))
        self.cps = np.zeros(max(voc_size, 1), dtype=np.uint32)
        if self.simple:
            for i in range(1, voc_size):
                self.cps[i] = ord(i2c[i])

    def strings(self, chars, lens):
        """chars int32 [B,S], lens int32 [B] -> list of B strings (index 0, the empty string, is dropped)."""
        B = chars.shape[0]
        if not self.simple:
            return [''.join(self.i2c[int(v)] for v in chars[j, :lens[j]]) for j in range(B)]
        S = chars.shape[1]
        keep = (np.arange(S)[None, :] < np.asarray(lens)[:, None]) & (chars > 0)
        counts = keep.sum(axis=1)
        text = self.cps[chars[keep]].tobytes().decode('utf-32-le')
        ends = np.cumsum(counts)
        return [text[e - c:e] for c, e in zip(counts.tolist(), ends.tolist())]


def greedy_line_score(probs_row):
    """decode_sequence_greedy's score (seq2seq.py:1318, 1330, 1341, 1353): ``decoded_score = 0`` then
    ``decoded_score += -np.log(prob)`` with float32 terms.  Under the reference's pinned numpy (< 1.24, legacy
    promotion) ``int + np.float32`` is a float64, so the running sum is float64 from the first character on:
    float32 logarithms, added left to right in float64 (cumsum is strictly sequential), divided by the number
    of characters."""
    p = np.asarray(probs_row, dtype=np.float32)
    if not len(p):
        return 0
    return np.cumsum(-np.log(p), dtype=np.float64)[-1] / len(p)


class Prefetcher:
    """Iterate over ``source`` in a worker thread, ``depth`` items ahead.  Exceptions raised by the
    source re-raise in the consumer at the position they occurred."""
    _END = object()

    def __init__(self, source, depth=2):
        self.q = queue.Queue(maxsize=depth)
        self.stop = threading.Event()
        self.thread = threading.Thread(target=self._run, args=(source,), daemon=True)
        self.thread.start()

    def _put(self, item):
        while not self.stop.is_set():
            try:
                self.q.put(item, timeout=0.1)
                return True
            except queue.Full:
                continue
        return False

    def _run(self, source):
        try:
            for item in source:
                if not self._put((item, None)):
                    return
        except BaseException as err:    # noqa: B902 — handed to the consumer
            self._put((None, err))
            return
        self._put((self._END, None))

    def __iter__(self):
        try:
            while True:
                item, err = self.q.get()
                if err is not None:
                    raise err
                if item is self._END:
                    return
                yield item
        finally:
            self.stop.set()


def chunked(iterable, n):
    """Lists of up to n consecutive items."""
    chunk = []
    for item in iterable:
        chunk.append(item)
        if len(chunk) == n:
            yield chunk
            chunk = []
    if chunk:
        yield chunk


# ----------------------------------------------------------------------------- line files -> reference batches

class LineFiles:
    """The line stream of ``Sequence2Sequence.gen_lines`` (ocrd_cor_asv_ann/lib/seq2seq.py:919-1018) as a reader
    object: records come out of ``records()`` one at a time (text files are streamed, pickles are loaded once),
    ``batches()`` groups them the way the reference's consumers expect.

    File formats: text files hold ``source<TAB>target`` lines (or plain lines when ``unsupervised``); ``*.pkl``
    files hold (source, target) pairs whose source is a string, a list of (char, prob) tuples or a confusion
    network (list of chunks of (string, prob) alternatives) and already ends in a newline.
    """

    SPLIT_RATIO = 0.2       # share of the lines that belongs to the validation generator (seq2seq.py:927)

    def __init__(self, filenames, split=None, train=False, unsupervised=False, charmap=None, reject=None):
        self.filenames = list(filenames)
        self.split = split if isinstance(split, np.ndarray) else None
        self.train, self.unsupervised = train, unsupervised
        self.table = str.maketrans(charmap) if charmap else None
        self.reject = reject            # callable(source, target) -> True to drop a training pair (bad OCR)
        self.soft = False               # the file being read carries confidences

    @staticmethod
    def _from_pickle(item):
        source, target = item
        if not source:
            text, conf = '', []
        elif type(source[0]) is tuple:                  # one (char, prob) per position
            conf = [p for _, p in source]
            text = ''.join(c for c, _ in source)
        else:                                           # confusion network: first alternative of every chunk
            conf = source
            text = ''.join(chunk[0][0] if chunk else '' for chunk in source)
        if not text.endswith('\n'):                     # degenerate entry: becomes an empty line
            text, conf = '\n', [[('\n', 1.0)]]
        return text, conf, target

    def _from_text(self, line):
        if self.unsupervised and '\t' not in line:
            return line, None, line
        source, target = line.split('\t')
        return source + '\n', None, target             # the file iterator left the newline on the target

    def records(self):
        """(source text, source confidences or None, target text, filename) per usable line, in file order."""
        for filename in self.filenames:
            self.soft = filename.endswith('.pkl')
            with open(filename, 'rb' if self.soft else 'r') as handle:
                items = pickle.load(handle) if self.soft else handle
                for line_no, item in enumerate(items):
                    if self.split is not None and (self.split[line_no] < self.SPLIT_RATIO) == self.train:
                        continue                        # this line belongs to the other generator
                    source, conf, target = self._from_pickle(item) if self.soft else self._from_text(item)
                    if self.unsupervised:
                        target = source
                    if self.table:
                        source, target = source.translate(self.table), target.translate(self.table)
                    source = unicodedata.normalize('NFC', source)
                    target = unicodedata.normalize('NFC', target)
                    if self.train and self.reject is not None and self.reject(source, target):
                        continue
                    yield source, conf, target, filename

    def batches(self, batch_size, pad_last):
        """4-tuples (sources, confidences or None, targets, filenames) of ``batch_size`` lines; the remainder is
        padded with empty lines when ``pad_last``, dropped otherwise (seq2seq.py:997-1017)."""
        src, cnf, tgt, names = [], [], [], []
        for source, conf, target, filename in self.records():
            src.append(source)
            tgt.append(target)
            names.append(filename)
            if self.soft:
                cnf.append(conf)
            if len(src) == batch_size:
                yield src, (cnf if self.soft else None), tgt, names
                src, cnf, tgt, names = [], [], [], []
        if src and pad_last:
            pad = batch_size - len(src)
            yield (src + [''] * pad, (cnf + [[] for _ in range(batch_size - len(cnf))]) if self.soft else None,
                   tgt + [''] * pad, names + [None] * pad)

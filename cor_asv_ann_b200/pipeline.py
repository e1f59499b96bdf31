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
import queue
import threading

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


def float32_mean_neglog(probs_row):
    """decode_sequence_greedy's score (seq2seq.py:1312, 1343, 1352): the float32 running sum of
    -log p in output order, divided by the number of characters.  cumsum adds strictly left to right
    in float32, i.e. exactly the reference's ``score += -np.log(p)`` chain."""
    p = np.asarray(probs_row, dtype=np.float32)
    if not len(p):
        return 0
    return np.cumsum(-np.log(p), dtype=np.float32)[-1] / len(p)


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

"""Oracle, part 3: hard re-alignment path from the soft attention matrix.

TEST INFRASTRUCTURE — see ``oracle/__init__.py``.

Restates ``_alignment2path`` of /root/reference/ocrd_cor_asv_ann/wrapper/transcode.py:279-349 (the
only consumer of the decoders' ``alignments``, called per line at :126).  PINNED: unlike the neural
part this function is plain numpy in the reference, so ``tests/ref_harness.py`` executes the
reference's own source here and ``tests/golden/ref_alignment2path.npz`` holds its outputs.

Semantics kept on purpose (they are what the wrapper ships):
  * forward scores are computed only for cell (0,0) and for cells whose alignment score exceeds
    ``min_score``; every other cell of the table stays 0 (:298-319);
  * the trace-back reads ``table[i-1, j]`` / ``table[i, j-1]`` with Python's negative indexing, so
    at i == 0 (j == 0) it looks at the LAST row (column) (:327-337);
  * strict ``>`` comparisons: on ties the diagonal step wins over both, and "left" (j-1) over "up";
  * start cell of the trace-back when the table is not square (:320-323), including the slice
    start that Python clips to 0;
  * ``dist`` adds ``1.0 - np.float32`` per visited cell: float64 under the reference's numpy<1.24,
    float32 under numpy>=2 (``oracle.decoders.PROMOTION``).
"""
import numpy as np

from . import decoders as _dec


def forward_table(alignment, i_max, j_max, min_score):
    """transcode.py:292-319.  Returns the (i_max, j_max) float32 table."""
    a = np.asarray([np.asarray(alignment[j], np.float32)[:i_max] for j in range(j_max)], np.float32)
    a = a.reshape(j_max, i_max)
    table = np.zeros((i_max, j_max), np.float32)
    if i_max == 0 or j_max == 0:
        return table
    if _dec.PROMOTION == "nep50":
        todo = a > np.float32(min_score)
    else:
        todo = a.astype(np.float64) > float(min_score)   # np.float32 > python float compares in float64 (numpy<1.24)
    todo[0, 0] = True                               # the walk starts at (0,0) unconditionally
    for j in range(j_max):
        for i in np.flatnonzero(todo[j]):
            best = np.float32(0)
            if i > 0:
                best = max(best, table[i - 1, j])
            if j > 0:
                best = max(best, table[i, j - 1])
            if i > 0 and j > 0:
                best = max(best, table[i - 1, j - 1])
            table[i, j] = a[j, i] + best
    return table


def alignment_path(alignment, i_max, j_max, min_score):
    """transcode.py:279-349.  Returns (path, dist): ``path`` int64 [i_max+1] with path[i] = output
    position of input position i (-1 where the reference's dict has no key), path[i_max] = j_max,
    path[0] = 0; ``dist`` = sum of (1 - score) over the visited cells."""
    assert i_max >= 1 and j_max >= 1
    a = [np.asarray(alignment[j], np.float32) for j in range(j_max)]
    table = forward_table(alignment, i_max, j_max, min_score)
    if i_max <= j_max:
        i = i_max - 1
    else:
        s = max(0, j_max - 2)
        i = j_max - 2 + int(np.argmax(table[s:, j_max - 1]))
    if j_max <= i_max:
        j = j_max - 1
    else:
        s = max(0, i_max - 2)
        j = i_max - 2 + int(np.argmax(table[i_max - 1, s:]))
    path = np.full(i_max + 1, -1, np.int64)
    path[i_max] = j_max
    nep50 = _dec.PROMOTION == "nep50"
    dist = np.float32(0) if nep50 else np.float64(0)
    while i >= 0 and j >= 0:
        if nep50:
            dist = np.float32(dist + (np.float32(1) - a[j][i]))
        else:
            dist = dist + (1.0 - np.float64(a[j][i]))
        path[i] = j
        up = table[(i - 1) % i_max, j]              # negative index wraps to the last row
        left = table[i, (j - 1) % j_max]
        diag = table[(i - 1) % i_max, (j - 1) % j_max]
        if up > left:
            if up > diag:
                i -= 1
            else:
                i -= 1
                j -= 1
        elif left > diag:
            j -= 1
        else:
            i -= 1
            j -= 1
    path[0] = 0
    return path, float(dist)

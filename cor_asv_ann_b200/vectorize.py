"""Input encoding for the engine: lines -> non-zeros of the reference's dense (B,T,V) array.

Mirrors the encoder half of ``vectorize_lines`` (ocrd_cor_asv_ann/lib/seq2seq.py:1020-1119):
unit vectors for plain strings, the confidence value for (char, prob) input, one entry per
alternative for confusion networks (later alternatives overwrite earlier ones at the same
index, like the dense assignment does), true zero for padding, index 0 for unmapped characters.
"""
import numpy as np

GAP = '\a'


_TABLES = {}


def _lookup_table(c2i):
    """Code point -> index as one dense array (-1 = unmapped), cached per mapping object.  Alphabets
    are a few hundred code points of the BMP; the table is capped at the full Unicode range."""
    key = (id(c2i), len(c2i))
    hit = _TABLES.get(key)
    if hit is not None and hit[0] is c2i:
        return hit[1]
    cps = [ord(c) for c in c2i if len(c) == 1]
    table = np.full(max(cps + [0]) + 2, -1, dtype=np.int32)      # last slot: everything beyond the alphabet
    for c, i in c2i.items():
        if len(c) == 1:
            table[ord(c)] = i
    if len(_TABLES) > 16:
        _TABLES.clear()
    _TABLES[key] = (c2i, table)
    return table


def strings_to_sparse(c2i, lines, logger=None):
    """Plain strings -> (idx int32 [B,T,1], None).  Vectorised: one pass over the joined text."""
    B = len(lines)
    lens = np.fromiter((len(s) for s in lines), dtype=np.int64, count=B)
    T = int(lens.max()) if B else 0
    idx = np.full((B, max(T, 1), 1), -1, dtype=np.int32)
    if T == 0:
        return idx[:, :0], None
    text = ''.join(lines)
    cps = np.frombuffer(text.encode('utf-32-le'), dtype=np.uint32)
    table = _lookup_table(c2i)
    mapped = table[np.minimum(cps, len(table) - 1)]
    hit = mapped >= 0
    if not hit.all():
        if logger is not None:
            for cp in np.unique(cps[~hit]):
                if chr(cp) != GAP:
                    logger.error('unmapped character "%s" in encoder input', chr(cp))
        mapped = np.where(hit, mapped, 0).astype(np.int32)
    if (lens == T).all():
        idx[:, :, 0] = mapped.reshape(B, T)
        return idx, None
    rows = np.repeat(np.arange(B), lens)
    cols = np.arange(len(cps)) - np.repeat(np.cumsum(lens) - lens, lens)
    idx[rows, cols, 0] = mapped
    return idx, None


def lines_to_sparse(c2i, lines, conf=None, logger=None):
    """General form.  Returns (idx int32 [B,T,k], val float32 [B,T,k] or None)."""
    if not conf:
        return strings_to_sparse(c2i, lines, logger)

    def lookup(ch, i, j):
        if ch in c2i:
            return c2i[ch]
        if ch != GAP and logger is not None:
            logger.error('unmapped character "%s" at encoder input sequence %d position %d', ch, i, j)
        return 0

    if type(conf[0][0]) is list:                    # confusion networks
        cells = []                                  # per line: list of per-position dicts idx -> prob
        for i, line in enumerate(conf):
            row = []
            for chunk in line:
                width = max(len(alt[0]) for alt in chunk) if chunk else 0
                start = len(row)
                row.extend({} for _ in range(width))
                for chars, p in chunk:
                    for k, ch in enumerate(chars):
                        row[start + k][lookup(ch, i, start + k)] = p
            cells.append(row)
    else:                                           # strings with per-character confidence
        cells = [[{lookup(ch, i, j): conf[i][j]} for j, ch in enumerate(line)] for i, line in enumerate(lines)]
    B = len(cells)
    T = max(len(r) for r in cells)
    k = max([len(c) for r in cells for c in r] + [1])
    idx = np.full((B, T, k), -1, dtype=np.int32)
    val = np.zeros((B, T, k), dtype=np.float32)
    for i, row in enumerate(cells):
        for j, cell in enumerate(row):
            for a, (v, p) in enumerate(cell.items()):
                idx[i, j, a] = v
                val[i, j, a] = p
    return idx, val


def dense_to_sparse(x):
    """Dense (B,T,V) rows (what the reference's decoders take) -> (idx, val)."""
    x = np.asarray(x)
    B, T, _ = x.shape
    nz = x != 0
    k = max(int(nz.sum(axis=2).max()) if x.size else 1, 1)
    idx = np.full((B, T, k), -1, dtype=np.int32)
    val = np.zeros((B, T, k), dtype=np.float32)
    b, t, v = np.nonzero(nz)
    if len(b):
        flat = b * T + t
        first = np.r_[True, flat[1:] != flat[:-1]]
        start = np.maximum.accumulate(np.where(first, np.arange(len(flat)), 0))
        slot = np.arange(len(flat)) - start
        idx[b, t, slot] = v
        val[b, t, slot] = x[b, t, v]
    if x.dtype.kind in 'ui' and (not len(b) or (val[idx >= 0] == 1).all()):
        return idx, None
    return idx, val


def rejection_table(idx, val):
    """Per (line, position): nanargmax of the dense source row, -1 where the row is all zero
    (``np.any(source_scores)`` / ``np.nanargmax(source_scores)``, seq2seq.py:1458-1462).
    First maximum wins, i.e. the lowest index among equal values."""
    v = np.ones(idx.shape, np.float32) if val is None else val
    live = (idx >= 0) & (v != 0)
    B, T, k = idx.shape
    # a dense row also holds zeros: a positive maximum beats them, a negative one does not
    key = np.where(live, v, -np.inf)
    best = key.max(axis=2)
    cand = np.where(live & (key == best[..., None]), idx, np.iinfo(np.int32).max)
    rej = cand.min(axis=2).astype(np.int32)
    rej[~live.any(axis=2)] = -1
    return rej

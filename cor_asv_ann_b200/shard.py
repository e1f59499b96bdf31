"""Multi-GPU sharding of the hot path: lines are independent once the reference's batches are fixed,
so ranks take disjoint blocks of WHOLE reference batches and there is no data-path collective
(SURVEY.md §8e).  A batch is never split across ranks because padding is not masked in the
reference: a line's result depends on the padded length of its batch
(``gen_lines`` forms 256-line batches in file order, seq2seq.py:997-1003).

One process per GPU (torchrun); ``torch.distributed`` is used only to return results to the caller
in input order (host objects), never inside the decode.
"""


def batch_ranges(n_lines, batch_size):
    """[(start, end)] of the reference's batches over n_lines lines in order."""
    return [(s, min(s + batch_size, n_lines)) for s in range(0, n_lines, batch_size)]


def partition_batches(n_lines, batch_size, world_size):
    """Contiguous blocks of whole batches per rank, sizes differing by at most one batch.
    Returns a list (one entry per rank) of lists of (start, end) line ranges."""
    batches = batch_ranges(n_lines, batch_size)
    base, extra = divmod(len(batches), world_size)
    out, pos = [], 0
    for r in range(world_size):
        n = base + (1 if r < extra else 0)
        out.append(batches[pos:pos + n])
        pos += n
    return out


def correct_lines_sharded(correct_batch, lines, conf=None, batch_size=256, rank=0, world_size=1, group=None):
    """Run ``correct_batch(lines, conf)`` (e.g. ``Sequence2Sequence.correct_lines`` with the mode
    arguments bound) on this rank's batches and return the 4 result lists for ALL lines in input order
    on every rank.  With world_size == 1 no process group is touched."""
    mine = partition_batches(len(lines), batch_size, world_size)[rank]
    local = []
    for s, e in mine:
        res = correct_batch(lines[s:e], conf[s:e] if conf else None)
        local.append((s, e, res))
    if world_size == 1:
        parts = [local]
    else:
        import torch.distributed as dist
        parts = [None] * world_size
        dist.all_gather_object(parts, local, group=group)
    out = ([None] * len(lines), [None] * len(lines), [None] * len(lines), [None] * len(lines))
    for part in parts:
        for s, e, res in part:
            for acc, vals in zip(out, res):
                acc[s:e] = vals
    return out

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


# ----------------------------------------------------------------------------- fixed jobs as arrays (strong scaling)

class JobResult:
    """Results of a whole job in input order, as arrays: chars int16 or int32 [n, S] (index 0 = nothing), lens int32 [n],
    scores float64 [n] and, for the beamed decoder, stats int64 [n, 8].  Strings are made on demand."""

    def __init__(self, chars, lens, scores, stats=None):
        self.chars, self.lens, self.scores, self.stats = chars, lens, scores, stats

    def __len__(self):
        return len(self.lens)

    def lines(self, i2c, voc_size):
        from .pipeline import StringTable
        return StringTable(i2c, voc_size).strings(self.chars, self.lens)


class LocalPart:
    """The job as one rank sees it: ``len`` is the size of the WHOLE job, but only this rank's lines [lo, lo + len(mine))
    exist here (e.g. its part of a shared file).  Supports exactly what decode_job does: slices inside that range."""

    def __init__(self, n_total, lo, mine):
        self.n_total, self.lo, self.mine = n_total, lo, mine

    def __len__(self):
        return self.n_total

    def __getitem__(self, k):
        assert isinstance(k, slice) and k.start >= self.lo and k.stop <= self.lo + len(self.mine), "line outside this rank's part"
        return self.mine[k.start - self.lo:k.stop - self.lo]


def engine_runner(s2s, fast=True, greedy=True):
    """run_group for decode_job on the CUDA engine of a configured ``Sequence2Sequence``: a group of reference
    batches goes through ONE engine call with device-resident results (cor_encode + cor_decode_greedy_batch, or
    cor_decode_beam_batch); returns device tensors (chars [B, 2T], lens [B], scores [B], stats or None)."""
    import torch
    assert not fast or greedy, "cannot decode in fast mode with beam search enabled"
    eng = s2s._get_engine()
    dev = eng.device

    def prepare(batches):
        """host side of a group (runs in the producer thread): index encoding, merge, pinned staging"""
        from .pipeline import merge_batches
        from .vectorize import lines_to_sparse
        encoded = [lines_to_sparse(s2s.mapping[0], lines, conf, s2s.logger) for lines, conf in batches]
        idx, val, row_T, _ = merge_batches(encoded)
        pin = lambda a: None if a is None else torch.from_numpy(a).pin_memory()
        return pin(idx), pin(val), pin(row_T)

    def run(staged):
        idx, val, row_T = (None if t is None else t.to(dev, non_blocking=True) for t in staged)
        if greedy:
            eng.encode(idx, val, row_T, want_outputs=False)
            chars, _, scores, lens, _ = eng.greedy()
            return chars, lens, scores, None
        bp = eng.beam_params(batch_size=min(int(s2s.batch_size), 256), beam_width_in=s2s.beam_width_in,
                             beam_threshold_in=s2s.beam_threshold_in, beam_width_out=s2s.beam_width_out,
                             rejection_threshold=s2s.rejection_threshold)
        chars, _, scores, lens, _, stats = eng.beam_device(idx, val, row_T, bp, budget_bytes=s2s.beam_budget_bytes)
        return chars, lens, scores, stats

    return prepare, run, dev


def decode_job(runner, lines, conf=None, batch_size=256, superbatch=64, rank=0, world_size=1, group=None):
    """Decode the fixed job ``lines`` (a sequence every rank can slice; a rank only touches its own part) on
    ``world_size`` ranks and return a ``JobResult`` for ALL lines in input order on rank 0 (None elsewhere).

    Ranks take contiguous blocks of whole reference batches (``partition_batches``); inside a rank, ``superbatch``
    reference batches share one engine call and the host side of the next group (index encoding, pinned staging)
    runs in a producer thread while the GPU decodes the current one.  Results stay arrays end to end: each rank
    fills one result tensor, and ONE gather per array brings them to rank 0 -- no Python object crosses ranks.
    ``runner`` = (prepare, run, device) as returned by ``engine_runner``."""
    import torch
    from .pipeline import Prefetcher, chunked
    prepare, run, dev = runner
    mine = partition_batches(len(lines), batch_size, world_size)[rank]
    n_local = sum(e - s for s, e in mine)
    per_rank = [sum(e - s for s, e in part) for part in partition_batches(len(lines), batch_size, world_size)]
    n_max = max(per_rank + [1])
    t_local = max([len(l) for s, e in mine for l in lines[s:e]] + [1])
    t_all = torch.tensor([t_local], dtype=torch.int64, device=dev)
    if world_size > 1:
        import torch.distributed as dist
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX, group=group)
    S = 2 * int(t_all.item())
    chars = torch.zeros((n_max, S), dtype=torch.int32, device=dev)
    lens = torch.zeros((n_max,), dtype=torch.int32, device=dev)
    scores = torch.zeros((n_max,), dtype=torch.float64, device=dev)
    stats = None

    def groups():
        for chunk in chunked(mine, max(int(superbatch), 1)):
            yield prepare([(lines[s:e], conf[s:e] if conf else None) for s, e in chunk])

    off = 0
    for staged in Prefetcher(groups(), depth=2):
        c, l, sc, st = run(staged)
        b = c.shape[0]
        chars[off:off + b, :c.shape[1]] = c
        lens[off:off + b] = l
        scores[off:off + b] = sc
        if st is not None:
            if stats is None:
                stats = torch.zeros((n_max, st.shape[1]), dtype=torch.int64, device=dev)
            stats[off:off + b] = st
        off += b
    assert off == n_local

    def to_host(t):
        """device -> pinned host memory (a pageable .cpu() of the 1 M-line result runs at a fraction of the link rate)"""
        if not t.is_cuda:
            return t
        host = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        host.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return host

    def to_root(t):
        if world_size == 1:
            return to_host(t[:n_local])
        import torch.distributed as dist
        wire = t.view(torch.uint8) if t.dtype == torch.int16 else t       # NCCL has no 16-bit integer type: bytes on the wire
        parts = [torch.empty_like(wire) for _ in range(world_size)] if rank == 0 else None
        dist.gather(wire, parts, dst=0, group=group)
        if rank != 0:
            return None
        return to_host(torch.cat([p[:n] for p, n in zip(parts, per_rank)]).view(t.dtype))

    # character indices travel as int16 when the alphabet allows (decided over all ranks: the gather needs one dtype):
    # half the bytes through the gather and the D2H copy
    if chars.is_cuda:
        cmax = chars.max().reshape(1).to(torch.int64) if chars.numel() else torch.zeros(1, dtype=torch.int64, device=dev)
        if world_size > 1:
            import torch.distributed as dist
            dist.all_reduce(cmax, op=dist.ReduceOp.MAX, group=group)
        if int(cmax.item()) < 32768:
            chars = chars.to(torch.int16)

    out = [to_root(t) for t in (chars, lens, scores)]
    st_out = to_root(stats) if stats is not None else None
    if rank != 0:
        return None
    return JobResult(out[0].numpy(), out[1].numpy(), out[2].numpy(), None if st_out is None else st_out.numpy())

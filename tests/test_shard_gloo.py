"""N>1 host logic on CPU: world-size-2 gloo run of the batch sharding.  Each rank decodes
its own whole reference batches (here with the oracle standing in for the GPU engine) and every rank
ends up with all results in input order, identical to the single-process run."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from cor_asv_ann_b200.shard import partition_batches, correct_lines_sharded, batch_ranges


def test_partition_keeps_batches_whole():
    for n, bs, world in [(1000, 256, 2), (256, 256, 8), (10240, 256, 8), (5, 2, 3), (0, 256, 2)]:
        parts = partition_batches(n, bs, world)
        assert len(parts) == world
        flat = [r for p in parts for r in p]
        assert flat == batch_ranges(n, bs)                       # nothing lost, order kept
        assert all(s % bs == 0 for s, _ in flat)                 # reference batch boundaries preserved
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lines, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import decoders as od
    from tests.common import make_model
    topo, w, mapping = make_model("c1")
    fn = lambda ls, cf: od.correct_lines(topo, w, mapping, ls, conf=cf, fast=True)
    out = correct_lines_sharded(fn, lines, None, batch_size=3, rank=rank, world_size=world)
    ret[rank] = (out[0], [float(x) for x in out[2]])
    dist.destroy_process_group()


def test_two_ranks_gloo_equal_single_process():
    from oracle import decoders as od
    from tests.common import make_model, synthetic_lines
    topo, w, mapping = make_model("c1")
    lines = synthetic_lines(mapping, 10, 9, seed=5, min_length=3)
    fn = lambda ls, cf: od.correct_lines(topo, w, mapping, ls, conf=cf, fast=True)
    single = correct_lines_sharded(fn, lines, None, batch_size=3)
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), lines, ret), nprocs=2, join=True)
    for rank in (0, 1):
        text, scores = ret[rank]
        assert text == single[0]
        assert np.allclose(scores, single[2], atol=1e-6)
    # batches differ in padded length, so decoding all 10 lines as ONE batch gives other numbers:
    # the reason a batch is never split across ranks
    whole = od.correct_lines(topo, w, mapping, lines, fast=True)
    assert not np.allclose(whole[2], single[2], atol=1e-6)


# ----------------------------------------------------------------------------- decode_job (arrays, one gather per array)

def _oracle_runner(topo, w, mapping):
    """decode_job's (prepare, run, device) with the oracle standing in for the engine, on CPU tensors."""
    import torch
    from oracle import decoders as od

    def prepare(batches):
        return batches

    def run(batches):
        texts, scores = [], []
        for lines, conf in batches:
            r = od.correct_lines(topo, w, mapping, list(lines), conf=conf, fast=True)
            texts += r[0]
            scores += [float(s) for s in r[2]]
        S = 2 * max(max(len(l) for l in b[0]) for b in batches)
        chars = torch.zeros((len(texts), S), dtype=torch.int32)
        for j, t in enumerate(texts):
            chars[j, :len(t)] = torch.tensor([mapping[0][c] for c in t], dtype=torch.int32)
        return chars, torch.tensor([len(t) for t in texts], dtype=torch.int32), torch.tensor(scores, dtype=torch.float64), None

    return prepare, run, torch.device("cpu")


def _job_worker(rank, world, port, lines, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cor_asv_ann_b200.shard import decode_job
    from tests.common import make_model
    topo, w, mapping = make_model("c1")
    res = decode_job(_oracle_runner(topo, w, mapping), lines, batch_size=3, superbatch=2, rank=rank, world_size=world)
    ret[rank] = None if res is None else (res.lines(mapping[1], topo.voc_size), res.scores.tolist())
    dist.destroy_process_group()


def test_decode_job_two_ranks_equals_one():
    from cor_asv_ann_b200.shard import decode_job
    from oracle import decoders as od
    from tests.common import make_model, synthetic_lines
    topo, w, mapping = make_model("c1")
    lines = synthetic_lines(mapping, 11, 9, seed=6, min_length=3)
    single = decode_job(_oracle_runner(topo, w, mapping), lines, batch_size=3, superbatch=2)
    want = [od.correct_lines(topo, w, mapping, lines[s:s + 3], fast=True) for s in range(0, 11, 3)]
    assert single.lines(mapping[1], topo.voc_size) == [t for r in want for t in r[0]]
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_job_worker, args=(2, _free_port(), lines, ret), nprocs=2, join=True)
    assert ret[1] is None                                   # results live on rank 0 only
    text, scores = ret[0]
    assert text == single.lines(mapping[1], topo.voc_size)  # input order
    assert np.allclose(scores, single.scores, atol=1e-12)

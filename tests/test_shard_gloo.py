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

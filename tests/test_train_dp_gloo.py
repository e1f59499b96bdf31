"""Data-parallel training protocol on CPU (world size 2, gloo), with the training ORACLE standing in for the CUDA step:
every rank computes the gradient of the SUM of its rows' weighted cross-entropies -- rows cut from a batch that was
vectorised to the batch's common padded lengths first (the model does not mask padding) -- the gradient vectors and the
(loss sum, target steps) pairs are all-reduced, and the result equals the single-process gradient of the whole batch.
This is the protocol cor_asv_ann_b200.trainer.Trainer.train_on_batch / training.run_epoch follow on GPUs over NCCL."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import model as om
from oracle import train as ot
from tests.common import synthetic_alphabet, synthetic_lines


def _setup():
    topo = om.Topology(width=16, depth=2, voc_size=30, residual_connections=True)
    w = om.init_weights(topo, 9, "U")
    mapping = synthetic_alphabet(30)
    src = synthetic_lines(mapping, 6, 8, seed=3, min_length=2)
    tgt = synthetic_lines(mapping, 6, 9, seed=4, min_length=2)
    return topo, w, mapping, src, tgt


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _flat(grads):
    return np.concatenate([g.reshape(-1) for k in sorted(grads) for g in grads[k]])


def _worker(rank, world, port, ret):
    from cor_asv_ann_b200.training import _share
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    topo, w, mapping, src, tgt = _setup()
    x, di, do, wt = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)         # the WHOLE batch: common padded lengths
    lo, hi = _share(len(src), rank, world)
    _, xent, grads = ot.loss_and_grads(topo, w, x[lo:hi], di[lo:hi], do[lo:hi], with_regulariser=False)
    n = float(wt[lo:hi].sum())
    g = torch.from_numpy(_flat(grads) * n)                                       # gradient of the SUM over this rank's rows
    pair = torch.tensor([xent * n, n], dtype=torch.float64)
    dist.all_reduce(g)
    dist.all_reduce(pair)
    ret[rank] = (g.numpy() / pair[1].item(), pair[0].item() / pair[1].item())
    dist.destroy_process_group()


def test_two_ranks_equal_single_process_gradient():
    topo, w, mapping, src, tgt = _setup()
    x, di, do, wt = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
    _, xent, grads = ot.loss_and_grads(topo, w, x, di, do, with_regulariser=False)
    ret = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    want = _flat(grads)
    for rank in (0, 1):
        g, loss = ret[rank]
        assert abs(loss - xent) < 1e-12
        assert np.abs(g - want).max() < 1e-12
    # cutting the batch BEFORE vectorising (each share padded to its own lengths) gives another gradient: why the rows
    # are cut from the vectorised batch
    short = [i for i in range(3)]
    x2, di2, do2, _ = ot.teacher_arrays(mapping, topo.voc_size, [src[i] for i in short], [tgt[i] for i in short])
    if x2.shape[1] != x.shape[1]:
        _, xent2, _ = ot.loss_and_grads(topo, w, x2, di2, do2, with_regulariser=False)
        _, xent1, _ = ot.loss_and_grads(topo, w, x[:3], di[:3], do[:3], with_regulariser=False)
        assert abs(xent1 - xent2) > 1e-9

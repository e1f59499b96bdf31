"""Worker of tests/test_gpu_train.py::test_data_parallel_step_two_gpus: launched by torch.distributed.run with one
process per GPU (NCCL).  Every rank runs Trainer.train_on_batch on its share of one batch; rank 0 also runs the whole
batch on a second, group-less trainer and compares the updated weights."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from cor_asv_ann_b200.trainer import Trainer, teacher_indices
    from cor_asv_ann_b200.training import _share
    from cor_asv_ann_b200 import synthetic
    W, d, V = 64, 2, 48
    topo = dict(width=W, depth=d, voc_size=V, residual_connections=True, window_width=5)
    w = synthetic.random_weights(synthetic.layer_shapes(W, d, V), W, d, np.random.default_rng(1), "U")
    c2i, i2c = synthetic.synthetic_alphabet(V)
    lines = synthetic.indices_to_lines(synthetic.line_indices(16, 12, V, 5, min_length=4), i2c)
    tgts = synthetic.indices_to_lines(synthetic.line_indices(16, 13, V, 6, min_length=4), i2c)
    s_idx, d_in, d_out = teacher_indices(c2i, lines, tgts)
    tr = Trainer(topo, w, device=local)
    lo, hi = _share(len(lines), rank, world)
    losses = [tr.train_on_batch(s_idx[lo:hi], None, d_in[lo:hi], d_out[lo:hi]) for _ in range(3)]
    ok = 1
    if rank == 0:
        ref = Trainer(topo, w, device=local)
        ref.all_reduce = lambda x, n: (x, n)            # single process: the whole batch, no collective
        ref_losses = [ref.train_on_batch(s_idx, None, d_in, d_out) for _ in range(3)]
        err = float((tr.params - ref.params).abs().max())
        dl = max(abs(a - b) for a, b in zip(losses, ref_losses))
        print("data parallel over %d GPUs vs one process: max |dw| %.3g, max |dloss| %.3g" % (world, err, dl))
        ok = int(err < 2e-5 and dl < 1e-5)
    flag = torch.tensor([ok], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag[0]) == 1 else 1)


if __name__ == "__main__":
    main()

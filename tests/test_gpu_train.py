"""Row T: the CUDA training step (csrc/train.cu through cor_asv_ann_b200.trainer.Trainer) against the training oracle
(oracle/train.py, torch float64 autograd): loss, every gradient, and the Adam(clipnorm) update.

Tolerance: the GPU step is fp32 end to end (cuBLAS SGEMM without TF32, fp32 transcendentals); against float64 autograd
its gradients agree to ~1e-5 of the largest entry of each tensor -- asserted at 2e-4 relative to that entry, plus 1e-6."""
import numpy as np
import pytest
import torch

from oracle import model as om
from oracle import train as ot
from tests.common import make_model, synthetic_lines, topo_dict

pytestmark = pytest.mark.gpu


def _topo(width, depth, V, resid=False):
    return om.Topology(width=width, depth=depth, voc_size=V, residual_connections=resid)


def _case(width, depth, V, resid, n, length, seed, masks):
    topo = _topo(width, depth, V, resid)
    w = om.init_weights(topo, seed, "U")
    from tests.common import synthetic_alphabet
    mapping = synthetic_alphabet(V)
    src = synthetic_lines(mapping, n, length, seed=seed, min_length=3)
    tgt = synthetic_lines(mapping, n, length + 2, seed=seed + 1, min_length=3)
    rng = np.random.default_rng(seed)
    mk = None
    if masks:
        D = 2 * width if depth == 1 else width
        draw = lambda shape: (rng.random(shape) < 0.8).astype(np.float32) / np.float32(0.8)
        mk = {"enc": [draw(2 * width if k == 0 else width) for k in range(depth - 1)], "dec": [draw(width) for _ in range(depth - 1)],
              "cell": draw((n, width + D))}
    return topo, w, mapping, src, tgt, mk


def _compare(topo, w, mapping, src, tgt, masks):
    from cor_asv_ann_b200.trainer import Trainer, teacher_indices
    x, di, do, _ = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
    loss, xent, grads = ot.loss_and_grads(topo, w, x, di, do, masks)
    tr = Trainer(topo_dict(topo), w, device=0)
    s_idx, d_in, d_out = teacher_indices(mapping[0], src, tgt)
    ce_sum, n = tr.forward_backward(s_idx, None, d_in, d_out, masks)
    assert n == float((d_out >= 0).sum())
    assert abs(ce_sum / n - xent) < 1e-5 * max(1.0, abs(xent)), (ce_sum / n, xent)
    reg, gnorm = tr.apply(n, lr=0.0)              # scale + regulariser gradient, no update
    assert abs(xent + reg - loss) < 1e-5 * max(1.0, abs(loss))
    got = tr.get_gradients()
    worst = 0.0
    for layer, arrs in grads.items():
        for i, g in enumerate(arrs):
            scale = np.abs(g).max()
            err = np.abs(got[layer][i] - g).max()
            if scale > 1e-8:
                worst = max(worst, err / scale)
            assert err < 2e-4 * scale + 1e-6, (layer, i, err, scale)
    ref_norm = np.sqrt(sum(float((g.astype(np.float64) ** 2).sum()) for v in grads.values() for g in v))
    assert abs(gnorm - ref_norm) < 1e-4 * ref_norm
    return tr, grads, worst


@pytest.mark.parametrize("width,depth,V,resid,masks", [(32, 1, 40, False, False), (32, 2, 40, False, False), (32, 2, 40, True, True),
                                                      (48, 3, 50, True, True), (64, 4, 40, True, False)])
def test_loss_and_gradients(width, depth, V, resid, masks):
    topo, w, mapping, src, tgt, mk = _case(width, depth, V, resid, 5, 9, 11 + depth, masks)
    _, _, worst = _compare(topo, w, mapping, src, tgt, mk)
    print("depth %d width %d: worst gradient error relative to the tensor's largest entry %.3g" % (depth, width, worst))


def test_adam_clipnorm_update_and_second_step():
    from cor_asv_ann_b200.trainer import teacher_indices
    topo, w, mapping, src, tgt, mk = _case(32, 2, 40, True, 6, 8, 5, True)
    tr, grads, _ = _compare(topo, w, mapping, src, tgt, mk)
    s_idx, d_in, d_out = teacher_indices(mapping[0], src, tgt)
    state = {}
    ref_w = w
    for step in range(3):
        x, di, do, _ = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
        _, _, g = ot.loss_and_grads(topo, ref_w, x, di, do, mk)
        ref_w, _ = ot.adam_step(ref_w, g, state)
        tr.train_on_batch(s_idx, None, d_in, d_out, mk)
        got = tr.get_weights()
        for layer, arrs in ref_w.items():
            for i, a in enumerate(arrs):
                assert np.abs(got[layer][i] - a).max() < 2e-5, (step, layer, i)
    assert tr.last["grad_norm"] > 0 and tr.last["target_steps"] == float((d_out >= 0).sum())


def test_confidence_input_and_large_gradient_clip():
    """Soft (confidence) source input, and a batch whose gradient norm exceeds the clip: the update direction is scaled."""
    from cor_asv_ann_b200.trainer import Trainer, teacher_indices
    topo, w, mapping, src, tgt, _ = _case(32, 2, 40, False, 4, 7, 21, False)
    w = {k: [a * np.float32(3.0) if k.startswith("decoder") and a.ndim == 2 else a for a in v] for k, v in w.items()}
    s_idx, d_in, d_out = teacher_indices(mapping[0], src, tgt)
    rng = np.random.default_rng(2)
    val = rng.uniform(0.3, 1.0, size=s_idx.shape).astype(np.float32)
    x, di, do, _ = ot.teacher_arrays(mapping, topo.voc_size, src, tgt)
    xs = x.astype(np.float64) * np.where(s_idx >= 0, val, 0.0).max(axis=2, keepdims=True)
    loss, xent, grads = ot.loss_and_grads(topo, w, xs, di, do)
    tr = Trainer(topo_dict(topo), w, device=0)
    ce, n = tr.forward_backward(s_idx, val, d_in, d_out)
    assert abs(ce / n - xent) < 1e-5 * max(1.0, xent)
    tr.apply(n, lr=0.0)
    got = tr.get_gradients()
    for layer, arrs in grads.items():
        for i, g in enumerate(arrs):
            assert np.abs(got[layer][i] - g).max() < 2e-4 * np.abs(g).max() + 1e-6, (layer, i)


def test_train_loop_learns_and_model_decodes(tmp_path):
    """Sequence2Sequence.train on a small copy task (target = source): the validation loss falls over the epochs, the
    object ends up trained (status 2), per-epoch checkpoints were written, and the trained weights decode."""
    import os
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    rng = np.random.default_rng(0)
    alphabet = "abcdefgh"
    lines = [''.join(rng.choice(list(alphabet), size=int(rng.integers(4, 10)))) for _ in range(320)]
    path = str(tmp_path / "train.txt")
    with open(path, "w") as f:
        f.write(''.join("%s\t%s\n" % (l, l) for l in lines))
    s = Sequence2Sequence()
    s.width, s.depth, s.batch_size, s.epochs, s.dropout = 64, 2, 32, 4, 0.0
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        s.train([path])
    finally:
        os.chdir(cwd)
    vl = s.history["val_loss"]
    print("val_loss per epoch:", ["%.3f" % v for v in vl])
    assert s.status == 2 and len(vl) >= 2 and vl[-1] < vl[0]
    assert any(n.startswith("model.ckpt.weights-") for n in os.listdir(tmp_path))
    out = s.correct_lines([l + "\n" for l in lines[:4]], None, fast=True, greedy=True)
    assert len(out[0]) == 4


def test_data_parallel_step_two_gpus():
    """NCCL all-reduce of the flat gradient vector: two ranks on halves of a batch == one process on the whole batch."""
    import os, subprocess, sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(root, "tests", "dp_train_worker.py")], capture_output=True, text=True, timeout=600)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0

"""Host side of row T (BASELINE.json config 5): data-parallel training steps on the CUDA trainer (csrc/train.cu).

One process per GPU.  A step is what ``encoder_decoder_model.train_on_batch`` does in the reference
(ocrd_cor_asv_ann/lib/keras_train.py:195 on the graph of seq2seq.py:316-390 compiled by :491-497): teacher-forced
forward, temporal-weighted categorical cross-entropy + embedding regulariser, ``Adam(clipnorm=5)``.  With a process
group, the flat gradient vector and the (loss sum, target steps) pair are all-reduced -- NCCL over NVLink on GPUs --
before every rank applies the identical update: the result equals one single-GPU step on the union of the batches.

PyTorch is plumbing (device vectors, streams, ``torch.distributed``); the arithmetic is in libcor_b200.so.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import CorConfig, CorTrainBatch, check


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def teacher_indices(c2i, sources, targets):
    """vectorize_lines (seq2seq.py:1020-1117) for plain strings as index arrays: (src_idx [B,Ts,1], dec_in [B,Td],
    dec_out [B,Td]) with -1 for "all-zero vector" (start symbol, padding); Td = longest target + 1."""
    B, Ts, Td = len(sources), max(map(len, sources)), max(map(len, targets)) + 1
    src = np.full((B, Ts, 1), -1, np.int32)
    di = np.full((B, Td), -1, np.int32)
    do = np.full((B, Td), -1, np.int32)
    for i, (s, t) in enumerate(zip(sources, targets)):
        src[i, :len(s), 0] = [c2i.get(ch, 0) for ch in s]
        idx = [c2i.get(ch, 0) for ch in t]
        di[i, 1:len(t) + 1] = idx
        do[i, :len(t)] = idx
    return src, di, do


class Trainer:
    """Flat fp32 parameter / gradient / Adam-moment vectors of one model on one GPU and the step on them.

    ``topology``: dict like Engine's (width, depth, voc_size, residual_connections, window_width);
    ``weights``: {keras_layer: [ndarray, ...]} in Keras order and layout (LSTM format)."""

    def __init__(self, topology, weights=None, device=0, group=None):
        if not torch.cuda.is_available():
            raise RuntimeError("cor_asv_ann_b200 needs a CUDA device; there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device("cuda", device)
        self.group = group
        self.cfg = CorConfig(
            width=int(topology["width"]), depth=int(topology["depth"]), voc_size=int(topology["voc_size"]),
            residual_connections=int(bool(topology.get("residual_connections", False))),
            deep_bidirectional_encoder=int(bool(topology.get("deep_bidirectional_encoder", False))),
            bridge_dense=int(bool(topology.get("bridge_dense", False))), lm_predict=0,
            window_width=int(topology.get("window_width", 5)), eos_index=1, gemm_mode=0)
        self.W, self.d, self.V = self.cfg.width, self.cfg.depth, self.cfg.voc_size
        self.D = 2 * self.W if self.d == 1 else self.W
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self.lib.cor_trainer_create(C.byref(self.cfg), device, C.byref(h)))
        self.handle = h
        n = C.c_int64()
        check(self.lib.cor_trainer_param_count(self.handle, C.byref(n)))
        self.n_params = n.value
        z = lambda: torch.zeros(self.n_params, dtype=torch.float32, device=self.device)
        self.params, self.grads, self._m, self._v = z(), z(), z(), z()
        check(self.lib.cor_trainer_bind(self.handle, _ptr(self.params), _ptr(self.grads), _ptr(self._m), _ptr(self._v)))
        self.last = {}
        if weights is not None:
            self.set_weights(weights)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.cor_trainer_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def layer_shapes(self):
        from .synthetic import layer_shapes
        return layer_shapes(self.W, self.d, self.V)

    def _slot(self, layer, i):
        off, cnt = C.c_int64(), C.c_int64()
        check(self.lib.cor_trainer_slot(self.handle, layer.encode(), i, C.byref(off), C.byref(cnt)))
        return off.value, cnt.value

    def set_weights(self, weights):
        host = torch.zeros(self.n_params, dtype=torch.float32)
        for layer, shapes in self.layer_shapes().items():
            for i, shape in enumerate(shapes):
                a = np.ascontiguousarray(weights[layer][i], np.float32)
                assert tuple(a.shape) == tuple(shape), (layer, i, a.shape, shape)
                off, cnt = self._slot(layer, i)
                assert cnt == a.size, (layer, i, cnt, a.size)
                host[off:off + cnt] = torch.from_numpy(a.reshape(-1))
        self.params.copy_(host)

    def _unflatten(self, flat):
        host = flat.detach().cpu().numpy()
        out = {}
        for layer, shapes in self.layer_shapes().items():
            out[layer] = []
            for i, shape in enumerate(shapes):
                off, cnt = self._slot(layer, i)
                out[layer].append(host[off:off + cnt].reshape(shape).copy())
        return out

    def get_weights(self):
        return self._unflatten(self.params)

    def get_gradients(self):
        return self._unflatten(self.grads)

    def launch_count(self, reset=False):
        return int(self.lib.cor_trainer_launch_count(self.handle, int(reset)))

    def reset_optimizer(self):
        check(self.lib.cor_trainer_reset_optimizer(self.handle))

    # ------------------------------------------------------------------ one step
    def forward_backward(self, src_idx, src_val, dec_in, dec_out, masks=None):
        """This rank's batch: src_idx int32 [B,Ts,k], src_val float32 or None, dec_in / dec_out int32 [B,Td] (numpy or
        tensors).  Leaves the gradient of the SUM of weighted cross-entropies in ``self.grads``; returns
        (cross-entropy sum, target steps)."""
        dev = self.device
        to = lambda a, dt: None if a is None else torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a, dtype=dt).to(dev).contiguous()
        src_idx, src_val = to(src_idx, torch.int32), to(src_val, torch.float32)
        dec_in, dec_out = to(dec_in, torch.int32), to(dec_out, torch.int32)
        B, Ts, k = src_idx.shape
        Td = dec_in.shape[1]
        masks = masks or {}
        me = md = mc = None
        if masks.get("enc") is not None and self.d > 1:
            me = torch.ones((self.d - 1, 2 * self.W), dtype=torch.float32)
            for n, m in enumerate(masks["enc"][:self.d - 1]):
                if m is not None:
                    me[n, :len(m)] = torch.as_tensor(np.asarray(m, np.float32))
            me = me.to(dev)
        if masks.get("dec") is not None and self.d > 1:
            md = torch.ones((self.d - 1, self.W), dtype=torch.float32)
            for n, m in enumerate(masks["dec"][:self.d - 1]):
                if m is not None:
                    md[n] = torch.as_tensor(np.asarray(m, np.float32))
            md = md.to(dev)
        if masks.get("cell") is not None:
            mc = to(np.asarray(masks["cell"], np.float32), torch.float32)
            assert tuple(mc.shape) == (B, self.W + self.D)
        self._keep = (src_idx, src_val, dec_in, dec_out, me, md, mc)
        bt = CorTrainBatch(B, Ts, k, Td, _ptr(src_idx), _ptr(src_val), _ptr(dec_in), _ptr(dec_out), _ptr(me), _ptr(md), _ptr(mc))
        out = (C.c_double * 2)()
        with torch.cuda.device(dev):
            check(self.lib.cor_train_forward_backward(self.handle, C.byref(bt), out, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out[0], out[1]

    def all_reduce(self, xent_sum, n_steps):
        """Sum the gradient vector and the loss pair over the ranks of ``self.group`` (no-op without one)."""
        import torch.distributed as dist
        if self.group is None and not (dist.is_available() and dist.is_initialized()):
            return xent_sum, n_steps
        dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=self.group)
        pair = torch.tensor([xent_sum, n_steps], dtype=torch.float64, device=self.device)
        dist.all_reduce(pair, op=dist.ReduceOp.SUM, group=self.group)
        return float(pair[0]), float(pair[1])

    def apply(self, n_steps, lr=1e-3, clipnorm=5.0, with_regulariser=True):
        """gradient / n_steps + regulariser gradient, global-norm clip, Adam.  Returns (regulariser value, gradient norm)."""
        out = (C.c_double * 2)()
        with torch.cuda.device(self.device):
            check(self.lib.cor_train_apply(self.handle, float(n_steps), int(with_regulariser), float(lr), float(clipnorm), out,
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out[0], out[1]

    def train_on_batch(self, src_idx, src_val, dec_in, dec_out, masks=None, lr=1e-3, clipnorm=5.0, with_regulariser=True):
        """One data-parallel step; returns the loss Keras would report (global mean cross-entropy + regulariser)."""
        xent, n = self.forward_backward(src_idx, src_val, dec_in, dec_out, masks)
        xent, n = self.all_reduce(xent, n)
        reg, gnorm = self.apply(n, lr, clipnorm, with_regulariser)
        self.last = {"xent": xent / max(n, 1.0), "regulariser": reg, "grad_norm": gnorm, "target_steps": n}
        return xent / max(n, 1.0) + (reg if with_regulariser else 0.0)

    def dropout_masks(self, rng, rate, n_rows):
        """Masks as the training graph samples them (seq2seq.py:293-298, 345, 363-367): one per feature between layers
        (time- and batch-constant), one per row and feature on the attention cell's input; scaled by 1 / keep."""
        if not rate:
            return None
        keep = 1.0 - rate
        draw = lambda shape: (rng.random(shape) < keep).astype(np.float32) / np.float32(keep)
        return {"enc": [draw(2 * self.W if n == 0 else self.W) for n in range(self.d - 1)],
                "dec": [draw(self.W) for n in range(self.d - 1)], "cell": draw((n_rows, self.W + self.D))}

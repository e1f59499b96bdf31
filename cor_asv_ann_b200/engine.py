"""Host-side handle on the CUDA engine: owns the model handle and a torch-allocated workspace.

PyTorch is plumbing here (device memory, pinned host memory, streams); all arithmetic runs
in libcor_b200.so.
"""
import ctypes as C
import numpy as np
import torch

from . import _lib
from ._lib import CorConfig, CorLines, CorBeamParams, check


BEAM_STATS = 8     # int64 statistics per line (COR_BEAM_STATS in include/cor_b200.h)
STAT_NAMES = ("iterations", "hypothesis_steps", "finals", "open_left", "score_ties", "near_ties", "pop_checksum", "fringe_peak")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class Engine:
    """One model on one GPU.

    ``topology``: dict with width, depth, voc_size and the optional switches
    residual_connections, deep_bidirectional_encoder, bridge_dense, lm_predict, window_width.
    ``weights``: {keras_layer_name: [ndarray, ...]} in Keras order and (in, out) layout
    (SURVEY.md A.5), LSTM layers already in LSTM (not CuDNN) format.
    """

    def __init__(self, topology, weights, device=0, eos_index=1, gemm_mode=0):
        if not torch.cuda.is_available():
            raise RuntimeError("cor_asv_ann_b200 needs a CUDA device; there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device("cuda", device)
        self.cfg = CorConfig(
            width=int(topology["width"]), depth=int(topology["depth"]), voc_size=int(topology["voc_size"]),
            residual_connections=int(bool(topology.get("residual_connections", False))),
            deep_bidirectional_encoder=int(bool(topology.get("deep_bidirectional_encoder", False))),
            bridge_dense=int(bool(topology.get("bridge_dense", False))),
            lm_predict=int(bool(topology.get("lm_predict", False))),
            window_width=int(topology.get("window_width", 5)),
            eos_index=int(eos_index), gemm_mode=int(gemm_mode))
        self.W, self.d, self.V = self.cfg.width, self.cfg.depth, self.cfg.voc_size
        self.D = 2 * self.W if (self.d == 1 or self.cfg.deep_bidirectional_encoder) else self.W
        self.lm = bool(self.cfg.lm_predict)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self.lib.cor_model_create(C.byref(self.cfg), device, C.byref(h)))
            self.handle = h
            for layer, arrs in weights.items():
                for i, a in enumerate(arrs):
                    a = np.ascontiguousarray(a, dtype=np.float32)
                    shape = (C.c_int64 * a.ndim)(*a.shape)
                    check(self.lib.cor_model_set_weight(self.handle, layer.encode(), i,
                                                        a.ctypes.data_as(C.c_void_p), shape, a.ndim))
            check(self.lib.cor_model_finalize(self.handle))
        self._ws = None
        self._geom = None
        self.last_ties = None      # int32 [B]: near-tie argmaxes per line of the last greedy call (device or pinned host tensor)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.cor_model_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ workspace
    def workspace_bytes(self, n_rows, t_max, k_alt, want_align=False):
        n = C.c_size_t()
        check(self.lib.cor_workspace_bytes(self.handle, n_rows, t_max, k_alt, int(want_align), C.byref(n)))
        return n.value

    def _workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws

    def release_workspace(self):
        """Give the workspace back to the allocator (a process that holds several engines, one per model, calls this
        before switching: a beam workspace is sized by its budget, tens of GB)."""
        self._ws = None
        self._geom = None
        torch.cuda.empty_cache()

    @staticmethod
    def _stream():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    # ------------------------------------------------------------------ encoder
    def encode(self, idx, val, row_T, want_align=False, want_outputs=True):
        """idx int32 [B,T,k] (device), val float32 [B,T,k] or None, row_T int32 [B].
        Returns (enc_out [T,B,D], states [2d,B,W]); the encoded lines stay in the workspace.
        want_outputs=False skips the two copies (the decoders read the workspace, not these): enc_out alone is
        T*B*D*4 bytes -- 2.7 GB for 16 384 lines -- and returns (None, None)."""
        B, T, k = idx.shape
        ws = self._workspace(self.workspace_bytes(B, T, k, want_align))
        lines = CorLines(B, T, k, _ptr(row_T), _ptr(idx), _ptr(val))
        enc_out = torch.empty((T, B, self.D), dtype=torch.float32, device=self.device) if want_outputs else None
        states = torch.empty((2 * self.d, B, self.W), dtype=torch.float32, device=self.device) if want_outputs else None
        with torch.cuda.device(self.device):
            check(self.lib.cor_encode(self.handle, C.byref(lines), int(want_align), _ptr(ws), ws.numel(),
                                      _ptr(enc_out), _ptr(states), self._stream()))
        self._geom = (B, T, k, bool(want_align))
        return enc_out, states

    def set_encoding(self, enc_out=None, states=None):
        """Install (or, without arguments, remove) encoder results to be imported instead of running the encoder:
        enc_out [T,B,D], states [2d,B,W] float32 device tensors (cor_set_encoding); kept alive by this object."""
        if enc_out is None:
            check(self.lib.cor_set_encoding(self.handle, 0, 0, None, None))
            self._preset = None
            return
        enc_out = enc_out.to(self.device, torch.float32).contiguous()
        states = states.to(self.device, torch.float32).contiguous()
        T, B, D = enc_out.shape
        assert D == self.D and tuple(states.shape) == (2 * self.d, B, self.W), (enc_out.shape, states.shape)
        self._preset = (enc_out, states)
        check(self.lib.cor_set_encoding(self.handle, B, T, _ptr(enc_out), _ptr(states)))

    # ------------------------------------------------------------------ one decoder step
    def decode_step(self, p_in, h_in, c_in, a_in, row_line=None):
        """p_in [R,V], h_in/c_in [d,R,W], a_in [R,T] -> (p, lm or None, h, c, a)."""
        R = p_in.shape[0]
        T = self._geom[1]
        dev = self.device
        p_out = torch.empty((R, self.V), dtype=torch.float32, device=dev)
        lm_out = torch.empty((R, self.V), dtype=torch.float32, device=dev) if self.lm else None
        h_out = torch.empty((self.d, R, self.W), dtype=torch.float32, device=dev)
        c_out = torch.empty((self.d, R, self.W), dtype=torch.float32, device=dev)
        a_out = torch.empty((R, T), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cor_decode_step(self.handle, _ptr(self._ws), self._ws.numel(), R, _ptr(row_line),
                                           _ptr(p_in.contiguous()), _ptr(h_in.contiguous()), _ptr(c_in.contiguous()),
                                           _ptr(a_in.contiguous()), _ptr(p_out), _ptr(lm_out), _ptr(h_out),
                                           _ptr(c_out), _ptr(a_out), self._stream()))
        return p_out, lm_out, h_out, c_out, a_out

    # ------------------------------------------------------------------ fast / batch greedy
    def greedy(self):
        """decode_batch_greedy over the lines of the last encode(); device tensors."""
        B, T, _, want_align = self._geom
        dev = self.device
        chars = torch.empty((B, 2 * T), dtype=torch.int32, device=dev)
        probs = torch.empty((B, 2 * T), dtype=torch.float32, device=dev)
        scores = torch.empty((B,), dtype=torch.float64, device=dev)
        lens = torch.empty((B,), dtype=torch.int32, device=dev)
        align = torch.empty((2 * T, B, T), dtype=torch.float32, device=dev) if want_align else None
        ties = torch.empty((B,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cor_decode_greedy_batch(self.handle, _ptr(self._ws), self._ws.numel(), _ptr(chars),
                                                   _ptr(probs), _ptr(scores), _ptr(lens), _ptr(ties), _ptr(align), self._stream()))
        self.last_ties = ties
        return chars, probs, scores, lens, align

    def correct_greedy_host(self, idx, val, row_T, want_align=False, out=None):
        """Host arrays in, host arrays out (copies inside the call).
        idx int32 [B,T,k], val float32 or None, row_T int32 [B]: numpy or pinned torch CPU tensors."""
        idx, val, row_T = _as_cpu(idx, torch.int32), _as_cpu(val, torch.float32), _as_cpu(row_T, torch.int32)
        B, T, k = idx.shape
        ws = self._workspace(self.workspace_bytes(B, T, k, want_align))
        if out is None:
            out = self.host_outputs(B, T, want_align)
        chars, probs, scores, lens, align = out
        ties = self.last_ties
        if ties is None or ties.is_cuda or ties.numel() != B:
            ties = torch.empty((B,), dtype=torch.int32, pin_memory=True)
        lines = CorLines(B, T, k, _ptr(row_T), _ptr(idx), _ptr(val))
        with torch.cuda.device(self.device):
            check(self.lib.cor_correct_greedy_host(self.handle, C.byref(lines), _ptr(ws), ws.numel(), _ptr(chars),
                                                   _ptr(probs), _ptr(scores), _ptr(lens), _ptr(ties), _ptr(align), self._stream()))
        self.last_ties = ties
        return out

    @staticmethod
    def host_outputs(B, T, want_align=False):
        pin = torch.cuda.is_available()
        return (torch.empty((B, 2 * T), dtype=torch.int32, pin_memory=pin),
                torch.empty((B, 2 * T), dtype=torch.float32, pin_memory=pin),
                torch.empty((B,), dtype=torch.float64, pin_memory=pin),
                torch.empty((B,), dtype=torch.int32, pin_memory=pin),
                torch.empty((2 * T, B, T), dtype=torch.float32, pin_memory=pin) if want_align else None)

    # ------------------------------------------------------------------ beam
    def beam_params(self, batch_size=256, beam_width_in=15, beam_threshold_in=0.2, beam_width_out=16,
                    rejection_threshold=0.3, n_best=1):
        return CorBeamParams(int(batch_size), int(beam_width_in), int(beam_width_out), int(n_best),
                             float(beam_threshold_in), float(rejection_threshold or 0.0))

    def correct_beam_host(self, idx, val, row_T, bp, want_align=False, budget_bytes=32 << 30, out=None):
        idx, val, row_T = _as_cpu(idx, torch.int32), _as_cpu(val, torch.float32), _as_cpu(row_T, torch.int32)
        B, T, k = idx.shape
        n = C.c_size_t()
        check(self.lib.cor_beam_workspace_bytes(self.handle, B, T, k, C.byref(bp), int(budget_bytes), C.byref(n)))
        ws = self._workspace(n.value)
        if out is None:
            pin = True
            N = max(1, int(bp.n_best))
            slots = (B,) if N == 1 else (B, N)       # n_best > 1: N result slots per line, best first (length 0 = unused)
            out = (torch.empty(slots + (2 * T,), dtype=torch.int32, pin_memory=pin),
                   torch.empty(slots + (2 * T,), dtype=torch.float32, pin_memory=pin),
                   torch.empty(slots, dtype=torch.float64, pin_memory=pin),
                   torch.empty(slots, dtype=torch.int32, pin_memory=pin),
                   torch.empty(slots + (2 * T, T), dtype=torch.float32, pin_memory=pin) if want_align else None,
                   torch.empty((B, BEAM_STATS), dtype=torch.int64, pin_memory=pin))
        chars, probs, scores, lens, align, stats = out
        lines = CorLines(B, T, k, _ptr(row_T), _ptr(idx), _ptr(val))
        with torch.cuda.device(self.device):
            check(self.lib.cor_correct_beam_host(self.handle, C.byref(lines), C.byref(bp), _ptr(ws), ws.numel(),
                                                 _ptr(chars), _ptr(probs), _ptr(scores), _ptr(lens), _ptr(align),
                                                 _ptr(stats), self._stream()))
        return out

    def beam_workspace(self, B, T, k, bp, budget_bytes):
        n = C.c_size_t()
        check(self.lib.cor_beam_workspace_bytes(self.handle, B, T, k, C.byref(bp), int(budget_bytes), C.byref(n)))
        return self._workspace(n.value)

    def beam_device(self, idx, val, row_T, bp, budget_bytes=32 << 30, out=None):
        """Beam search with the lines already on the device; results stay on the device.
        Returns (chars [B,2T], probs, scores, lens, None, stats [B,8]) as device tensors."""
        B, T, k = idx.shape
        ws = self.beam_workspace(B, T, k, bp, budget_bytes)
        dev = self.device
        if out is None:
            out = (torch.empty((B, 2 * T), dtype=torch.int32, device=dev), torch.empty((B, 2 * T), dtype=torch.float32, device=dev),
                   torch.empty((B,), dtype=torch.float64, device=dev), torch.empty((B,), dtype=torch.int32, device=dev), None,
                   torch.empty((B, BEAM_STATS), dtype=torch.int64, device=dev))
        chars, probs, scores, lens, align, stats = out
        lines = CorLines(B, T, k, _ptr(row_T), _ptr(idx), _ptr(val))
        with torch.cuda.device(dev):
            check(self.lib.cor_decode_beam_batch(self.handle, C.byref(lines), C.byref(bp), _ptr(ws), ws.numel(),
                                                 _ptr(chars), _ptr(probs), _ptr(scores), _ptr(lens), _ptr(align),
                                                 _ptr(stats), self._stream()))
        return out

    # ------------------------------------------------------------------ Viterbi re-alignment (row N1)
    def alignment_paths(self, align, layout, in_len, out_len, min_score=None):
        """Hard path per line from a DEVICE alignment tensor (wrapper/transcode.py:279-349).
        layout "greedy": align [S,B,T] (greedy()/cor_decode_greedy_batch); "beam": align [B,S,T].
        in_len/out_len int32 [B] (device or host); out_len < 0 = identity matrix of that size.
        Returns (path int32 [B,T+1] with -1 for "no key", dist float64 [B]) as device tensors."""
        assert align.is_cuda and align.dtype == torch.float32 and align.is_contiguous()
        if layout == "greedy":
            S, B, T = align.shape
            stride_step, stride_row = B * T, T
        else:
            B, S, T = align.shape
            stride_step, stride_row = T, S * T
        dev = self.device
        in_len = torch.as_tensor(in_len, dtype=torch.int32).to(dev).contiguous()
        out_len = torch.as_tensor(out_len, dtype=torch.int32).to(dev).contiguous()
        path = torch.empty((B, T + 1), dtype=torch.int32, device=dev)
        dist = torch.empty((B,), dtype=torch.float64, device=dev)
        ms = 1.0 / self.V if min_score is None else float(min_score)
        with torch.cuda.device(dev):
            check(self.lib.cor_alignment_paths(_ptr(align), stride_step, stride_row, B, T, S, _ptr(in_len),
                                               _ptr(out_len), ms, _ptr(path), T + 1, _ptr(dist), self._stream()))
        return path, dist

    def prof_enable(self, on=True):
        check(self.lib.cor_prof_enable(self.handle, int(on)))

    def prof_read(self):
        """(total ms, launches, algorithmic flops) of the LSTM-GEMM launches since prof_enable(True)."""
        ms, n, fl = C.c_double(), C.c_int64(), C.c_double()
        check(self.lib.cor_prof_read(self.handle, C.byref(ms), C.byref(n), C.byref(fl)))
        return ms.value, n.value, fl.value

    def launch_count(self, reset=False):
        """Kernels run since the last reset (nodes of the loop graphs included)."""
        if reset:       # both counters share their reset
            v = int(self.lib.cor_launch_count(self.handle, 0))
            self.lib.cor_host_launch_count(self.handle, 1)
            return v
        return int(self.lib.cor_launch_count(self.handle, 0))

    def host_launch_count(self):
        """Launch calls the host made since the last reset (kernel launches + graph launches)."""
        return int(self.lib.cor_host_launch_count(self.handle, 0))

    KERNEL_KINDS = ("tc_store_narrow", "tc_store_single", "tc_store_pair", "tc_lstm_narrow", "tc_lstm_single", "tc_lstm_pair",
                    "ffma_store", "ffma_lstm")

    def kernel_counters(self, reset=False):
        """GEMM launches of this handle by kernel variant (cor_debug_counters): which kernel a test really ran."""
        out = (C.c_int64 * 8)()
        check(self.lib.cor_debug_counters(self.handle, out, int(reset)))
        return dict(zip(self.KERNEL_KINDS, (int(v) for v in out)))


def _as_cpu(x, dtype):
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    assert x.device.type == "cpu"
    return x.to(dtype).contiguous()

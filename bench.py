#!/usr/bin/env python
"""bench.py — corrected chars/sec of the cor-asv-ann neural hot path on N B200s.

Workloads (BASELINE.json configs):
  greedy  C3: depth 2 / width 512 / V 512, fast batch-greedy decode of synthetic 80-char lines
  beam    C2: same model, beamed decode (beam_width_in 4, rejection 0.5), per-line A* semantics
A "step" is one pass of the hot path (encode + decode) over one super-batch of
`--lines` synthetic lines per GPU (whole 256-line reference batches, T = 81).  chars = input
characters excluding the newline (80 per line).  Lines shard across ranks with no data-path
collective (weak scaling); the only collective is the MAX of the elapsed time.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload greedy|beam] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, DEPTH, VOC, LINE_LEN, REF_BATCH = 512, 2, 512, 80, 256
BEAM = dict(batch_size=256, beam_width_in=4, beam_threshold_in=0.2, beam_width_out=16, rejection_threshold=0.5)


def flops_enc(T, W=WIDTH, d=DEPTH):
    D = 2 * W if d == 1 else W
    return T * W * W * (32 + 24 * (d >= 2) + 16 * max(d - 2, 0)) + 2 * T * D * W


def flops_step(W=WIDTH, d=DEPTH, V=VOC):
    D = 2 * W if d == 1 else W
    return 4 * V * W + 16 * W * W * (d - 1) + 2 * W * W + 8 * W * (W + D) + 8 * W * W


def build_model(seed=20251):
    from oracle.model import Topology, init_weights
    from tests.common import synthetic_alphabet
    topo = Topology(width=WIDTH, depth=DEPTH, voc_size=VOC)
    return topo, init_weights(topo, seed, "U"), synthetic_alphabet(VOC)


def make_indices(n_lines, seed):
    """Synthetic lines directly as index arrays [n, 81, 1]: 80 chars uniform over 2..V-1, then '\\n' (=1)."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(2, VOC, size=(n_lines, LINE_LEN + 1, 1), dtype=np.int32)
    idx[:, LINE_LEN, 0] = 1
    return idx


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.rows, self.proc, self.gpu = [], None, gpu

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_oracle_rate(workload, n_lines, seed=77):
    """The oracle (numpy restatement of the reference schedule) on the host cores: chars/s on n_lines."""
    from oracle import decoders as od
    topo, w, mapping = build_model()
    idx = make_indices(n_lines, seed)[:, :, 0]
    x = np.zeros((n_lines, LINE_LEN + 1, VOC), np.uint32)
    x[np.arange(n_lines)[:, None], np.arange(LINE_LEN + 1)[None, :], idx] = 1
    t0 = time.perf_counter()
    if workload == "greedy":
        od.decode_batch_greedy(topo, w, mapping, x, hoist_u=False)   # u recomputed per step like the reference
    else:
        from oracle.model import encode
        enc = encode(topo, w, x)
        for j in range(n_lines):
            next(od.decode_sequence_beam(topo, w, mapping, x[j], [e[j:j + 1] for e in enc], **BEAM))
    dt = time.perf_counter() - t0
    return LINE_LEN * n_lines / dt, dt


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    """--impl reference: the reference's CPU path.  Keras/TF is not installable here, so
    this times the oracle port on all host cores, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_threads()
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is meant to use every host thread
    from threadpoolctl import threadpool_limits
    threadpool_limits(limits=cores)
    total = args.steps + args.warmup
    probe_lines = 8 if args.workload == "greedy" else 4
    rate, _ = cpu_oracle_rate(args.workload, probe_lines)
    budget = 150.0 / max(total, 1)
    n = int(min(REF_BATCH, max(probe_lines, rate * budget / LINE_LEN)))
    for _ in range(args.warmup):
        cpu_oracle_rate(args.workload, n)
    t = []
    for s in range(args.steps):
        r, dt = cpu_oracle_rate(args.workload, n, seed=100 + s)
        t.append(dt)
    dt = sum(t)
    value = LINE_LEN * n * args.steps / dt
    sample = "%d lines x %d steps of the %s workload (numpy oracle, OpenBLAS threads)" % (n, args.steps, args.workload)
    print(json.dumps({
        "impl": "reference", "metric": "corrected chars/sec (%s)" % args.workload, "value": value, "unit": "chars/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, n),
        "cpu_baseline": {"value": value, "unit": "chars/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "chars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(args, lines_per_step, workload=None):
    workload = workload or args.workload
    return {"workload": ("C3 fast batch-greedy decode" if workload == "greedy" else
                         "C2 beamed decode (beam_width_in 4, rejection 0.5)") +
                        ", depth 2 width 512 V 512, synthetic 80-char lines, T=81, random-init weights (variant U)",
            "lines_per_step_per_gpu": lines_per_step, "reference_batch": REF_BATCH,
            "l2": "inputs differ every step; per-step activations (GBs) exceed the 126 MB L2",
            "parallelism": "lines sharded by 256-line reference batch, no data-path collective"}


def measure(args, eng, workload, rank, world, local, sample_clocks):
    """Time `workload` on this rank: device-resident pass (CUDA events) and end-to-end host call."""
    import torch
    import torch.distributed as dist
    L = (args.lines // REF_BATCH) * REF_BATCH
    if workload == "beam":
        L = (min(args.lines, args.beam_lines) // REF_BATCH) * REF_BATCH
    T = LINE_LEN + 1
    total_steps = args.steps + args.warmup
    seed0 = 1003 if workload == "greedy" else 1002          # config 3 / config 2
    # distinct lines for every step and rank, resident in HBM before the timed region
    dev_idx = [torch.as_tensor(make_indices(L, seed0 + 7919 * rank + 101 * s), device="cuda") for s in range(total_steps)]
    row_T = torch.full((L,), T, dtype=torch.int32, device="cuda")
    bp = eng.beam_params(**BEAM)
    budget = int(args.beam_budget_gb * (1 << 30))
    beam_out = None

    def device_step(s):
        nonlocal beam_out
        if workload == "greedy":
            eng.encode(dev_idx[s], None, row_T, want_outputs=False)     # the decoder reads the workspace; no 2.7 GB copy of enc_out
            return eng.greedy()
        beam_out = eng.beam_device(dev_idx[s], None, row_T, bp, budget_bytes=budget, out=beam_out)
        return beam_out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for s in range(args.warmup):
        device_step(s)
    barrier()
    sampler = ClockSampler(local)
    if sample_clocks:
        sampler.start()
    eng.launch_count(reset=True)
    eng.prof_enable(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for s in range(args.warmup, total_steps):
        out = device_step(s)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = eng.launch_count()
    k_ms, k_n, k_flops = eng.prof_read()
    eng.prof_enable(False)
    clocks = sampler.stop() if sample_clocks else None
    if workload == "greedy":
        dec_steps = 2 * T * L                            # the fast decoder always runs 2T steps for every line
        extra = {"decoded_chars_last_step": int(out[3].sum().item())}
    else:
        stats = out[5].cpu().numpy()
        dec_steps = int(stats[:, 1].sum())               # hypothesis-steps actually run (last step)
        lens = out[3].cpu().numpy()
        same = int((out[0][:, :LINE_LEN + 1].cpu().numpy() == dev_idx[total_steps - 1][:, :, 0].cpu().numpy()).all(axis=1).sum())
        extra = {"beam_iterations_per_line": float(stats[:, 0].mean()), "hypothesis_steps_per_line": dec_steps / L,
                 "lines_without_final": int((lens == 0).sum()), "lines_equal_to_input": same}

    # ---- end to end through the C-ABI host call: pinned host buffers, H2D and D2H inside
    host_idx = [torch.as_tensor(make_indices(L, seed0 + 4001 + 7919 * rank + 101 * s)).pin_memory() for s in range(args.steps)]
    host_T = torch.full((L,), T, dtype=torch.int32).pin_memory()
    if workload == "greedy":
        host_out = eng.host_outputs(L, T)
        call = lambda s: eng.correct_greedy_host(host_idx[s], None, host_T, out=host_out)
        api = "cor_correct_greedy_host (pinned host buffers)"
    else:
        host_out = None
        def call(s):
            nonlocal host_out
            host_out = eng.correct_beam_host(host_idx[s], None, host_T, bp, budget_bytes=budget, out=host_out)
        api = "cor_correct_beam_host (pinned host buffers)"
    call(0)
    barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        call(s)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    h2d = host_idx[0].numel() * 4 + host_T.numel() * 4
    d2h = sum(t.numel() * t.element_size() for t in host_out if t is not None)

    t_ms = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t_ms[0]), float(t_ms[1])
    chars = LINE_LEN * L * args.steps * world
    algo_flops_line = flops_enc(T) + dec_steps / L * flops_step()
    del dev_idx, host_idx
    torch.cuda.empty_cache()
    return {"L": L, "ms": ms, "value": chars / (ms * 1e-3), "e2e_value": chars / (e2e_ms * 1e-3), "e2e_ms": e2e_ms,
            "h2d": h2d, "d2h": d2h, "api": api, "launches": launches, "k_ms": k_ms, "k_n": k_n, "k_flops": k_flops,
            "clocks": clocks, "algo_flops_line": algo_flops_line, "extra": extra}


def measured_traffic(workload, rows):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/traffic.json), or None when no capture exists for this workload at this row count."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload)
    except (OSError, ValueError):
        return None, None
    if not t or t.get("rows") != rows:
        return None, None
    return t["bytes_per_launch"], t["source"]


def measure_page(args, eng, n_lines):
    """The small-batch regime (one OCR-D page per call, wrapper/transcode.py:111): fast greedy decode of
    `n_lines` lines through the host entry point; latency per call and per decoder timestep."""
    import torch
    T = LINE_LEN + 1
    idx = [torch.as_tensor(make_indices(n_lines, 2000 + s)).pin_memory() for s in range(4)]
    host_T = torch.full((n_lines,), T, dtype=torch.int32).pin_memory()
    out = eng.host_outputs(n_lines, T)
    for s in range(3):
        eng.correct_greedy_host(idx[s % 4], None, host_T, out=out)
    torch.cuda.synchronize()
    eng.launch_count(reset=True)
    reps = 10
    t0 = time.perf_counter()
    for s in range(reps):
        eng.correct_greedy_host(idx[s % 4], None, host_T, out=out)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    page = {"lines": n_lines, "ms_per_call": dt * 1e3, "us_per_decoder_timestep": dt * 1e6 / (2 * T + T),
            "timesteps_per_call": 3 * T, "chars_per_s": LINE_LEN * n_lines / dt,
            "gpu_launches_per_call": eng.launch_count() / reps,
            "api": "cor_correct_greedy_host (pinned host buffers), T encoder + 2T decoder timesteps"}
    # the OCR-D processor's default mode (fast_mode false, ocrd-tool.json:55-58) is the beamed decoder: same page,
    # config-2 beam parameters (random-init weights: T iterations of one hypothesis per line)
    bp = eng.beam_params(**BEAM)
    b_out = None
    for s in range(2):
        b_out = eng.correct_beam_host(idx[s % 4], None, host_T, bp, budget_bytes=2 << 30, out=b_out)
    torch.cuda.synchronize()
    eng.launch_count(reset=True)
    t0 = time.perf_counter()
    for s in range(reps):
        b_out = eng.correct_beam_host(idx[s % 4], None, host_T, bp, budget_bytes=2 << 30, out=b_out)
    torch.cuda.synchronize()
    db = (time.perf_counter() - t0) / reps
    page["beam"] = {"ms_per_call": db * 1e3, "chars_per_s": LINE_LEN * n_lines / db, "gpu_launches_per_call": eng.launch_count() / reps,
                    "api": "cor_correct_beam_host (pinned host buffers), persistent encoder recurrence + per-iteration beam kernels"}
    return page


def measure_predict(args, topo, w, mapping, device, n_lines):
    """Row N2, the call a user of the reference makes (scripts/proc.py:62): Sequence2Sequence.predict on a
    text file, fast mode, strings in -> strings + scores out; 64 reference batches per engine call with
    the next super-batch read and encoded in a worker thread, against the reference's one-batch loop."""
    import tempfile
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    s = Sequence2Sequence()
    s.width, s.depth, s.voc_size, s.mapping = topo.width, topo.depth, topo.voc_size, mapping
    s.device = device
    s.configure()
    for layer, arrs in w.items():
        s.encoder_decoder_model.get_layer(layer).set_weights(arrs)
    s.status = 2
    i2c = np.array([ord(mapping[1][i]) if mapping[1][i] else 0 for i in range(VOC)], dtype=np.uint32)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        def run(n, superbatch, seed):
            path = os.path.join(tmp, "lines_%d_%d.txt" % (n, superbatch))
            idx = make_indices(n, seed)[:, :, 0]
            with open(path, "w", encoding="utf-8") as f:
                f.write(i2c[idx].tobytes().decode("utf-32-le"))
            s.superbatch = superbatch
            t0 = time.perf_counter()
            n_out = sum(len(lines) for _, lines, _ in s.predict([path], fast=True, greedy=True))
            dt = time.perf_counter() - t0
            assert n_out == n
            return LINE_LEN * n / dt
        run(4 * REF_BATCH, 4, 3000)                       # warm-up (engine creation, workspaces)
        out["chars_per_s"] = run(n_lines, 64, 3001)
        out["chars_per_s_one_batch_per_call"] = run(min(n_lines, 8 * REF_BATCH), 1, 3002)
    out.update({"lines": n_lines, "superbatch": 64, "api": "Sequence2Sequence.predict(text file, fast=True), file read + encode + decode + strings"})
    # one OCR-D page through the class the wrapper calls (wrapper/transcode.py:111-115): strings in -> strings,
    # probabilities, scores, alignments out, fast greedy mode and the default beamed mode
    page = [i2c[make_indices(1, 5000 + j)[0, :, 0]].tobytes().decode("utf-32-le") for j in range(32)]
    s.rejection_threshold, s.beam_width_in = BEAM["rejection_threshold"], BEAM["beam_width_in"]
    for fast in (True, False):
        for _ in range(2):
            s.correct_lines(page, None, fast=fast, greedy=fast)
        t0 = time.perf_counter()
        for _ in range(5):
            res = s.correct_lines(page, None, fast=fast, greedy=fast)
        assert len(res[0]) == len(page)
        out["page32_correct_lines_ms_%s" % ("fast" if fast else "beam")] = (time.perf_counter() - t0) / 5 * 1e3
    return out


def bench_line(args, r, workload, world, peaks):
    peak = peaks.get("bf16_tflops_sustained", 1400.0)
    traffic, traffic_src = measured_traffic(workload, r["L"])
    peak_src = "MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks else "fallback 1400 (of fallback)"
    achieved = r["k_flops"] / (r["k_ms"] * 1e-3) / 1e12 if r["k_ms"] > 0 else 0.0
    line = {
        "metric": "corrected chars/sec (%s)" % workload, "value": r["value"], "unit": "chars/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"] / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16" if args.gemm_mode == 3 else "f32", "data": "synthetic", "config": workload_config(args, r["L"], workload),
        "e2e": {"value": r["e2e_value"], "unit": "chars/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                "ms_per_step": r["e2e_ms"] / args.steps, "api": r["api"]},
        "gpu_launches": r["launches"],
        "clocks": r["clocks"],
        "roofline": {"bound": "tensor", "kernel": "GEMM + fused LSTM-cell epilogue", "achieved": achieved,
                     "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "launches": r["k_n"], "kernel_ms_per_step": r["k_ms"] / args.steps,
                     "kernel_share_of_step": r["k_ms"] / r["ms"] if r["ms"] else None,
                     "precision": {1: "fp32 via FFMA", 3: "fp16 single pass: the reduced-precision VARIANT with the looser tolerance stated in DESIGN.md section 3, not the product default"}.get(args.gemm_mode, "fp32-grade (see DESIGN.md)")},
        "algorithmic_gflop_per_line": r["algo_flops_line"] / 1e9,
        "whole_step_tflops": r["algo_flops_line"] * r["L"] * args.steps * world / (r["ms"] * 1e-3) / 1e12,
    }
    line.update(r["extra"])
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=os.environ.get("COR_BENCH_WORKLOAD", "beam"), choices=["greedy", "beam"])
    ap.add_argument("--lines", type=int, default=int(os.environ.get("COR_BENCH_LINES", "16384")))
    ap.add_argument("--beam-lines", type=int, default=9472)  # 37 x 256 lines: the GEMM tile count is a whole number of waves on 74 CTA pairs
    ap.add_argument("--beam-budget-gb", type=float, default=72.0)
    ap.add_argument("--gemm-mode", type=int, default=0)
    ap.add_argument("--cpu-lines", type=int, default=None)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--page-lines", type=int, default=32, help="lines of the small-batch (one page) latency leg; 0 = skip")
    ap.add_argument("--page-only", action="store_true", help="run just the small-batch latency leg")
    ap.add_argument("--predict-lines", type=int, default=65536, help="lines of the Sequence2Sequence.predict leg; 0 = skip")
    ap.add_argument("--predict-only", action="store_true", help="run just the predict leg")
    ap.add_argument("--only", action="store_true", help="measure only --workload (skip the other decoder)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from cor_asv_ann_b200.engine import Engine
    from tests.common import topo_dict

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    # stdout carries exactly one JSON line: anything libraries print while we measure (NCCL's version
    # banner under NCCL_DEBUG=VERSION, for one) goes to stderr instead
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    topo, w, mapping = build_model()
    eng = Engine(topo_dict(topo), w, device=local, eos_index=1, gemm_mode=args.gemm_mode)
    if args.predict_only:
        r = measure_predict(args, topo, w, mapping, local, args.predict_lines)
        sys.stdout.flush(); os.dup2(saved_stdout, 1)
        print(json.dumps({"predict": r}))
        return
    if args.page_only:
        print_page = measure_page(args, eng, args.page_lines)
        sys.stdout.flush(); os.dup2(saved_stdout, 1)
        print(json.dumps({"page": print_page}))
        return
    other = "greedy" if args.workload == "beam" else "beam"
    main_r = measure(args, eng, args.workload, rank, world, local, sample_clocks=(rank == 0))
    other_r = None if args.only else measure(args, eng, other, rank, world, local, sample_clocks=(rank == 0))
    page_r = measure_page(args, eng, args.page_lines) if (rank == 0 and args.page_lines > 0 and not args.only) else None
    predict_r = measure_predict(args, topo, w, mapping, local, args.predict_lines) if (rank == 0 and world == 1 and args.predict_lines > 0 and not args.only) else None

    if world > 1:
        dist.barrier()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        line = bench_line(args, main_r, args.workload, world, peaks)
        if other_r is not None:
            o = bench_line(args, other_r, other, world, peaks)
            line["also"] = {k: o[k] for k in ("metric", "value", "unit", "ms_per_step", "config", "e2e", "gpu_launches",
                                               "clocks", "roofline", "algorithmic_gflop_per_line", "whole_step_tflops")}
        if page_r is not None:
            line["page"] = page_r
        if predict_r is not None:
            line["predict"] = predict_r
        if world == 1 and not args.no_cpu:
            cores = host_threads()
            from threadpoolctl import threadpool_limits
            threadpool_limits(limits=cores)
            n_cpu = args.cpu_lines or (256 if args.workload == "greedy" else 384)   # ~10-20 s of CPU work
            v, dt = cpu_oracle_rate(args.workload, n_cpu)
            line["cpu_baseline"] = {"value": v, "unit": "chars/s", "cores": cores, "kind": "port",
                                    "sample": "%d lines of the same workload, numpy oracle (Keras/TF unavailable), %.1f s"
                                              % (n_cpu, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- corrected chars/sec of the cor-asv-ann neural hot path on N B200s.

Workloads (BASELINE.json configs); every model is random-init of the named topology,
lines are synthetic (80 characters + newline, T = 81, seeded) unless stated:
  beam              C2 as configured: depth 2 / width 512 / V 512, beamed decode, beam_width_in 4, rejection 0.5
                    (weights variant U).  With random weights this is the DEGENERATE bracket: one hypothesis per
                    iteration, output == input.  It is the literal config and stays the headline `value`.
  greedy            C3: same model, fast batch-greedy decode (2T steps for every line).
  beam_competitive  C2 topology with PEAKED weights (variant P: a selective softmax, like a trained model's) and
                    the class-default beam parameters (beam_width_in 15, rejection 0.3): the search competes --
                    ~100 live hypotheses per line and iteration, finals found, early exit by beam_width_out.
  beam_stress       C2 bracket (ii): rejection_threshold 0 -- 256 hypotheses x 2T iterations per line, fringe at
                    its 2T*256 truncation, no final found (falls back to the input).
  c4                depth 3 / width 768 deep bidirectional encoder + residual + LM-rated beam (beam_width_in 15).
  c1                depth 1 / width 128 on the 55 lines of the reference's test pages, one call per page.
A "step" is one pass of the hot path (encode + decode) over one super-batch of lines per GPU (whole 256-line
reference batches); chars = input characters excluding the newline.  Lines shard across ranks with no data-path
collective (weak scaling: the `value`); `strong` is the literal fixed job of the config (C3: 1 000 000 lines,
C2: 100 000 lines) partitioned over the ranks by cor_asv_ann_b200.shard and gathered on rank 0 in input order.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload beam|greedy|...] [--impl reference]
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

LINE_LEN, REF_BATCH = 80, 256
C2 = dict(width=512, depth=2, voc_size=512)
C4 = dict(width=768, depth=3, voc_size=512, deep_bidirectional_encoder=True, residual_connections=True, lm_predict=True)
BEAM_C2 = dict(batch_size=256, beam_width_in=4, beam_threshold_in=0.2, beam_width_out=16, rejection_threshold=0.5)
BEAM_DEFAULT = dict(batch_size=256, beam_width_in=15, beam_threshold_in=0.2, beam_width_out=16, rejection_threshold=0.3)
WORKLOADS = {
    "beam": dict(topo=C2, variant="U", seed=1002, beam=BEAM_C2, lines=9472,
                 name="C2 beamed decode as configured (beam_width_in 4, rejection 0.5): the degenerate bracket, 1 hypothesis per iteration"),
    "greedy": dict(topo=C2, variant="U", seed=1003, beam=None, lines=16384, name="C3 fast batch-greedy decode"),
    "beam_competitive": dict(topo=C2, variant="P", seed=1012, beam=BEAM_DEFAULT, lines=2048,
                             name="C2 topology, peaked weights (variant P), class-default beam (beam_width_in 15, rejection 0.3): competing hypotheses"),
    "beam_stress": dict(topo=C2, variant="U", seed=1022, beam=dict(BEAM_C2, rejection_threshold=0.0), lines=256,
                        name="C2 stress bracket: rejection_threshold 0 (256 hypotheses x 2T iterations per line)"),
    "c4": dict(topo=C4, variant="U", seed=1004, beam=dict(BEAM_C2, beam_width_in=15), lines=1024,
               name="C4 depth 3 / width 768 deep-bidirectional + residual + LM-rated beam (beam_width_in 15, rejection 0.5)"),
}


def flops_enc(T, topo):
    W, d = topo["width"], topo["depth"]
    if topo.get("deep_bidirectional_encoder"):
        return T * W * W * (32 + 48 * (d - 1)) + 2 * T * (2 * W) * W
    D = 2 * W if d == 1 else W
    return T * W * W * (32 + 24 * (d >= 2) + 16 * max(d - 2, 0)) + 2 * T * D * W


def flops_step(topo):
    W, d, V = topo["width"], topo["depth"], topo["voc_size"]
    D = 2 * W if (d == 1 or topo.get("deep_bidirectional_encoder")) else W
    f = 4 * V * W + 16 * W * W * (d - 1) + 2 * W * W + 8 * W * (W + D) + 8 * W * W
    if topo.get("lm_predict"):
        f += 8 * W * (W + D) + 8 * W * W + 2 * V * W
    return f


def build_weights(topo, variant, seed=20251):
    """Seeded random-init weights of `topo` from the product's own generator (no oracle involved)."""
    from cor_asv_ann_b200 import synthetic
    shapes = synthetic.layer_shapes(topo["width"], topo["depth"], topo["voc_size"],
                                    topo.get("deep_bidirectional_encoder", False), topo.get("bridge_dense", False))
    return synthetic.random_weights(shapes, topo["width"], topo["depth"], np.random.default_rng(seed), variant)


def make_indices(n_lines, seed, voc=512):
    from cor_asv_ann_b200.synthetic import line_indices
    return line_indices(n_lines, LINE_LEN, voc, seed)


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.rows, self.proc, self.gpu, self.first = [], None, gpu, 0

    def mark(self):
        """Samples from here on count (the process is started before the warm-up: nvidia-smi's own start-up -- NVML
        initialisation takes the driver lock for a few hundred ms -- must not fall into the timed region)."""
        self.first = len(self.rows)

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
        for r in self.rows[self.first:]:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ----------------------------------------------------------------------------- the oracle legs (CPU baseline, parity)

def oracle_topology(topo):
    from oracle.model import Topology
    return Topology(**topo)


def cpu_oracle_rate(workload, n_lines, seed=77):
    """The oracle (numpy restatement of the reference schedule) on the host cores: chars/s on n_lines."""
    from oracle import decoders as od
    from cor_asv_ann_b200.synthetic import synthetic_alphabet
    spec = WORKLOADS[workload]
    topo, w = oracle_topology(spec["topo"]), build_weights(spec["topo"], spec["variant"])
    mapping = synthetic_alphabet(spec["topo"]["voc_size"])
    V = spec["topo"]["voc_size"]
    idx = make_indices(n_lines, seed, V)[:, :, 0]
    x = np.zeros((n_lines, LINE_LEN + 1, V), np.uint32)
    x[np.arange(n_lines)[:, None], np.arange(LINE_LEN + 1)[None, :], idx] = 1
    t0 = time.perf_counter()
    if spec["beam"] is None:
        od.decode_batch_greedy(topo, w, mapping, x, hoist_u=False)   # u recomputed per step like the reference
    else:
        from oracle.model import encode
        enc = encode(topo, w, x)
        for j in range(n_lines):
            next(od.decode_sequence_beam(topo, w, mapping, x[j], [e[j:j + 1] for e in enc], **spec["beam"]), None)
    dt = time.perf_counter() - t0
    return LINE_LEN * n_lines / dt, dt


def parity_check(workload, idx_rows, got, n_check):
    """Oracle on `n_check` lines of the LAST TIMED step's first reference batch against what the GPU returned for
    them: strings (and search statistics for the beam), per-character probabilities."""
    from oracle import decoders as od
    from oracle.model import encode
    from cor_asv_ann_b200.synthetic import synthetic_alphabet
    spec = WORKLOADS[workload]
    topo, w = oracle_topology(spec["topo"]), build_weights(spec["topo"], spec["variant"])
    V = spec["topo"]["voc_size"]
    mapping = synthetic_alphabet(V)
    batch = idx_rows[:REF_BATCH, :, 0]                 # the whole reference batch: its padded T is part of the input
    T = batch.shape[1]
    x = np.zeros((batch.shape[0], T, V), np.uint32)
    x[np.arange(batch.shape[0])[:, None], np.arange(T)[None, :], batch] = 1
    chars, probs, lens, stats = got
    t0 = time.perf_counter()
    same = ties = 0
    worst = 0.0
    diffs = []
    if spec["beam"] is None:
        n_check = batch.shape[0]
        text, r_probs, _, _ = od.decode_batch_greedy(topo, w, mapping, x)
        for j in range(n_check):
            g = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]])
            if g == text[j]:
                same += 1
                if lens[j]:
                    worst = max(worst, float(np.abs(probs[j, :lens[j]] - np.asarray(r_probs[j], np.float32)).max()))
    else:
        from tests.make_golden import stat_row
        enc = encode(topo, w, x)
        for j in range(n_check):
            st = {}
            r = next(od.decode_sequence_beam(topo, w, mapping, x[j], [e[j:j + 1] for e in enc], stats=st, **spec["beam"]), None)
            want = stat_row(st)
            ok = (r is None) == (lens[j] == 0) and list(stats[j, :5]) == want[:5] and stats[j, 6] == want[6]
            if ok and r is not None:
                ok = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) == r[0]
            if ok:
                same += 1
                if r is not None:
                    worst = max(worst, float(np.abs(probs[j, :lens[j]] - np.asarray(r[1], np.float32)).max()))
            else:
                if stats[j, 5] > 0 or want[5] > 0:
                    ties += 1
                got_s = ''.join(mapping[1][int(v)] for v in chars[j, :lens[j]]) if lens[j] > 0 else None
                diffs.append({"line": j, "string_equal": (r is None and got_s is None) or (r is not None and got_s == r[0]),
                              "gpu_stats[iterations,hyp_steps,finals,open_left,score_ties,near_ties]": [int(v) for v in stats[j, :6]],
                              "oracle_stats": [int(v) for v in want[:6]], "checksum_equal": bool(stats[j, 6] == want[6])})
    return {"parity_rows_checked": n_check, "rows_identical": same, "rows_differing_at_logged_near_tie": ties,
            "rows_differing_otherwise": n_check - same - ties, "max_abs_dp": worst, "differing_rows": diffs,
            "compared": "strings + per-character probabilities" + ("" if spec["beam"] is None else " + search statistics and expansion checksum"),
            "oracle_s": time.perf_counter() - t0}


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
    if args.workload == "train":
        n = args.cpu_lines or 64
        for _ in range(min(args.warmup, 1)):
            cpu_oracle_train_rate(n)
        t = [cpu_oracle_train_rate(n)[1] for _ in range(args.steps)]
        value = LINE_LEN * n * args.steps / sum(t)
        sample = "%d lines x %d steps of the C5 training step (torch float64 autograd oracle, %d threads)" % (n, args.steps, cores)
        print(json.dumps({"impl": "reference", "metric": "trained chars/sec (train)", "value": value, "unit": "chars/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(t) / args.steps, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": "C5 training step, %d lines per step" % n},
                          "cpu_baseline": {"value": value, "unit": "chars/s", "cores": cores, "kind": "port", "sample": sample},
                          "e2e": {"value": value, "unit": "chars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
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
        "config": workload_config(args.workload, n),
        "cpu_baseline": {"value": value, "unit": "chars/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "chars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(workload, lines_per_step):
    spec = WORKLOADS[workload]
    t = spec["topo"]
    return {"workload": "%s; depth %d width %d V %d, synthetic %d-char lines, T=%d, random-init weights (variant %s)"
                        % (spec["name"], t["depth"], t["width"], t["voc_size"], LINE_LEN, LINE_LEN + 1, spec["variant"]),
            "lines_per_step_per_gpu": lines_per_step, "reference_batch": REF_BATCH,
            "beam_params": spec["beam"],
            "l2": "inputs differ every step; per-step activations (GBs) exceed the 126 MB L2",
            "parallelism": "lines sharded by 256-line reference batch, no data-path collective"}


# ----------------------------------------------------------------------------- measurement

def measure(args, eng, workload, lines, rank, world, local, sample_clocks, roofline_pass=True, parity_lines=0):
    """Time `workload` on this rank: device-resident pass (CUDA events, nothing else inside the timed region), a
    separate pass with per-launch event bracketing for the roofline, and the end-to-end host call."""
    import torch
    import torch.distributed as dist
    spec = WORKLOADS[workload]
    L = max(REF_BATCH, (lines // REF_BATCH) * REF_BATCH)
    T = LINE_LEN + 1
    V = spec["topo"]["voc_size"]
    total_steps = args.steps + args.warmup
    seed0 = spec["seed"]                                    # config 3 seed 1003, config 2 seed 1002, ...
    # distinct lines for every step and rank, resident in HBM before the timed region
    host_steps = [make_indices(L, seed0 + 7919 * rank + 101 * s, V) for s in range(total_steps)]
    dev_idx = [torch.as_tensor(a, device="cuda") for a in host_steps]
    row_T = torch.full((L,), T, dtype=torch.int32, device="cuda")
    bp = eng.beam_params(**spec["beam"]) if spec["beam"] else None
    budget = int(args.beam_budget_gb * (1 << 30))
    beam_out = None

    def device_step(s):
        nonlocal beam_out
        if bp is None:
            eng.encode(dev_idx[s], None, row_T, want_outputs=False)     # the decoder reads the workspace; no 2.7 GB copy of enc_out
            return eng.greedy()
        beam_out = eng.beam_device(dev_idx[s], None, row_T, bp, budget_bytes=budget, out=beam_out)
        return beam_out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if sample_clocks:
        sampler.start()
    for s in range(args.warmup):
        device_step(s)
    barrier()
    sampler.mark()
    eng.launch_count(reset=True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev0.record()
    for s in range(args.warmup, total_steps):
        out = device_step(s)
        marks[s - args.warmup].record()         # an event record, no synchronisation: shows WHICH step an outlier sits in
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    ms_each = [round(a.elapsed_time(b), 2) for a, b in zip([ev0] + marks[:-1], marks)]
    launches, host_launches = eng.launch_count(), eng.host_launch_count()
    clocks = sampler.stop() if sample_clocks else None
    if bp is None:
        dec_steps = 2 * T * L                            # the fast decoder always runs 2T steps for every line
        extra = {"decoded_chars_last_step": int(out[3].sum().item()),
                 "near_tie_argmaxes_last_step": int(eng.last_ties.sum().item())}
        stats = None
    else:
        stats = out[5].cpu().numpy()
        dec_steps = int(stats[:, 1].sum())               # hypothesis-steps actually run (last step)
        lens = out[3].cpu().numpy()
        same = int((out[0][:, :LINE_LEN + 1].cpu().numpy() == dev_idx[total_steps - 1][:, :, 0].cpu().numpy()).all(axis=1).sum())
        extra = {"beam_iterations_per_line": float(stats[:, 0].mean()), "hypothesis_steps_per_line": dec_steps / L,
                 "hypotheses_per_line_and_iteration": dec_steps / max(int(stats[:, 0].sum()), 1),
                 "hypothesis_steps_per_s": dec_steps * args.steps * world / (ms * 1e-3) if ms else None,
                 "finals_per_line": float(stats[:, 2].mean()), "fringe_peak_mean": float(stats[:, 7].mean()),
                 "fringe_peak_max": int(stats[:, 7].max()), "near_tie_decisions_per_line": float(stats[:, 5].mean()),
                 "lines_without_final": int((lens == 0).sum()), "lines_equal_to_input": same}
    parity = None
    if parity_lines and rank == 0:
        got = (out[0].cpu().numpy(), out[1].cpu().numpy(), out[3].cpu().numpy(), stats)
        parity = parity_check(workload, host_steps[total_steps - 1], got, parity_lines)

    # ---- separate pass for the roofline: every LSTM-GEMM launch bracketed by two events (perturbs the step: not in `value`)
    k_ms = k_n = k_flops = 0
    prof_ms = None
    if roofline_pass:
        barrier()
        eng.prof_enable(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n_prof = min(args.steps, 2)
        for s in range(args.warmup, args.warmup + n_prof):
            device_step(s)
        e1.record()
        torch.cuda.synchronize()
        k_ms, k_n, k_flops = eng.prof_read()
        eng.prof_enable(False)
        prof_ms = e0.elapsed_time(e1)
        k_ms, k_n = k_ms / n_prof, k_n / n_prof
        k_flops /= n_prof
        prof_ms /= n_prof

    # ---- end to end through the C-ABI host call: pinned host buffers, H2D and D2H inside
    host_idx = [torch.as_tensor(make_indices(L, seed0 + 4001 + 7919 * rank + 101 * s, V)).pin_memory() for s in range(args.steps)]
    host_T = torch.full((L,), T, dtype=torch.int32).pin_memory()
    if bp is None:
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
    algo_flops_line = flops_enc(T, spec["topo"]) + dec_steps / L * flops_step(spec["topo"])
    del dev_idx, host_idx
    torch.cuda.empty_cache()
    return {"L": L, "ms": ms, "value": chars / (ms * 1e-3), "e2e_value": chars / (e2e_ms * 1e-3), "e2e_ms": e2e_ms,
            "h2d": h2d, "d2h": d2h, "api": api, "launches": launches, "host_launches": host_launches, "k_ms": k_ms, "k_n": k_n, "k_flops": k_flops,
            "prof_ms": prof_ms, "clocks": clocks, "algo_flops_line": algo_flops_line, "extra": extra, "parity": parity, "ms_each": ms_each}


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
            "gpu_launches_per_call": eng.launch_count() / reps, "host_launch_calls_per_call": eng.host_launch_count() / reps,
            "api": "cor_correct_greedy_host (pinned host buffers), T encoder + 2T decoder timesteps"}
    # the OCR-D processor's default mode (fast_mode false, ocrd-tool.json:55-58) is the beamed decoder: same page,
    # config-2 beam parameters (random-init weights: T iterations of one hypothesis per line)
    bp = eng.beam_params(**BEAM_C2)
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
                    "host_launch_calls_per_call": eng.host_launch_count() / reps,
                    "api": "cor_correct_beam_host (pinned host buffers)"}
    return page


def measure_early_exit(args, device, n_target=4096):
    """Fast decoder on lines that END (reference seq2seq.py:1255-1256 steps finished lines to 2T
    and ignores them).  Random-init decoders never emit a newline, so this leg uses weights of variant E (see
    cor_asv_ann_b200/synthetic.py) and, from a pool of candidate lines, those that an untimed first pass sees end within
    T of the 2T steps.  The loop graph compacts finished rows out of the GEMMs every 8 steps and stops when the last line
    has finished; the host loop (COR_NO_GRAPH=1) launches all 2T steps over all rows.  Results are identical
    (tests/test_gpu_loop_graph.py); only the time differs."""
    import torch
    from cor_asv_ann_b200.engine import Engine
    eng = Engine(dict(C2, window_width=5), build_weights(C2, "E"), device=device, eos_index=1, gemm_mode=args.gemm_mode)
    T = LINE_LEN + 1
    pool = make_indices(4 * n_target, 1031)
    keep = []
    for b0 in range(0, len(pool), n_target):
        idx = torch.as_tensor(pool[b0:b0 + n_target], device="cuda")
        eng.encode(idx, None, torch.full((len(idx),), T, dtype=torch.int32, device="cuda"), want_outputs=False)
        lens = eng.greedy()[3].cpu().numpy()
        keep.append(b0 + np.nonzero(lens < T)[0])
    keep = np.concatenate(keep)
    n = (min(len(keep), n_target) // REF_BATCH) * REF_BATCH
    if n == 0:
        return None
    idx = torch.as_tensor(pool[keep[:n]], device="cuda")
    row_T = torch.full((n,), T, dtype=torch.int32, device="cuda")
    out = {"workload": "C2 topology, weights of variant E (lines end), %d lines selected by a first pass: those that end within T of the 2T steps" % n, "lines": n}
    for name, env in (("loop_graph", None), ("host_loop_2T_steps", "1")):
        if env:
            os.environ["COR_NO_GRAPH"] = env
        for _ in range(2):
            eng.encode(idx, None, row_T, want_outputs=False); res = eng.greedy()
        torch.cuda.synchronize()
        eng.launch_count(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            eng.encode(idx, None, row_T, want_outputs=False); res = eng.greedy()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        lens = res[3].cpu().numpy()
        out[name] = {"ms_per_call": ms, "chars_per_s": LINE_LEN * n / (ms * 1e-3), "kernels_per_call": eng.launch_count() / 3,
                     "decoded_length_mean": float(lens.mean()), "decoded_length_max": int(lens.max())}
        os.environ.pop("COR_NO_GRAPH", None)
    out["speedup"] = out["host_loop_2T_steps"]["ms_per_call"] / out["loop_graph"]["ms_per_call"]
    eng.release_workspace()
    return out


def cpu_oracle_train_rate(n_lines):
    """The training oracle (torch float64 autograd, all host threads) on n_lines of the C5 workload: trained chars/s."""
    import torch
    from oracle import train as ot
    from oracle.model import Topology
    from cor_asv_ann_b200.synthetic import synthetic_alphabet, indices_to_lines
    torch.set_num_threads(host_threads())
    topo = Topology(**C2)
    w = build_weights(C2, "U")
    mapping = synthetic_alphabet(C2["voc_size"])
    lines = indices_to_lines(make_indices(n_lines, 1005), mapping[1])
    x, di, do, _ = ot.teacher_arrays(mapping, C2["voc_size"], lines, lines)
    t0 = time.perf_counter()
    _, _, g = ot.loss_and_grads(topo, w, x, di, do)
    ot.adam_step(w, g, {})
    dt = time.perf_counter() - t0
    return LINE_LEN * n_lines / dt, dt


def measure_train(args, local, rank, world, lines_per_gpu=64):
    """Config 5: training forward/backward + Adam(clipnorm 5), depth 2 / width 512 / V 512, 64 lines per GPU per step
    (source = target, 80 characters + newline: 81 encoder and 82 decoder steps), dropout 0.2 with fresh masks per step,
    data parallel: NCCL all-reduce of the flat fp32 gradient vector (13.4 M floats) and the loss pair every step."""
    import torch
    import torch.distributed as dist
    from cor_asv_ann_b200.trainer import Trainer, teacher_indices
    from cor_asv_ann_b200.synthetic import synthetic_alphabet, indices_to_lines
    mapping = synthetic_alphabet(C2["voc_size"])
    tr = Trainer(dict(C2, window_width=5), build_weights(C2, "U"), device=local)
    total = args.steps + args.warmup
    rng = np.random.default_rng(17 + rank)
    batches = []
    for s in range(total):
        lines = indices_to_lines(make_indices(lines_per_gpu, 1005 + 7919 * rank + 101 * s), mapping[1])
        batches.append(teacher_indices(mapping[0], lines, lines))
    dev = [tuple(torch.as_tensor(a, device="cuda") for a in b) for b in batches]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    losses = []
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for s in range(args.warmup):
        tr.train_on_batch(dev[s][0], None, dev[s][1], dev[s][2], tr.dropout_masks(rng, 0.2, lines_per_gpu))
    barrier()
    sampler.mark()
    tr.launch_count(reset=True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for s in range(args.warmup, total):
        losses.append(tr.train_on_batch(dev[s][0], None, dev[s][1], dev[s][2], tr.dropout_masks(rng, 0.2, lines_per_gpu)))
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    launches = tr.launch_count()
    # end to end: host (numpy) batches in, loss out -- the H2D copies of the index arrays inside
    barrier()
    t0 = time.perf_counter()
    for s in range(args.warmup, total):
        tr.train_on_batch(batches[s][0], None, batches[s][1], batches[s][2], tr.dropout_masks(rng, 0.2, lines_per_gpu))
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    t_ms = torch.tensor([ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t_ms[0]), float(t_ms[1])
    chars = LINE_LEN * lines_per_gpu * args.steps * world
    h2d = sum(a.nbytes for a in batches[0]) + (2 * 512 + 512 + lines_per_gpu * 1024) * 4
    flops_line = 3.0 * (flops_enc(LINE_LEN + 1, C2) + (LINE_LEN + 2) * flops_step(C2))
    return {"ms": ms, "value": chars / (ms * 1e-3), "e2e_value": chars / (e2e_ms * 1e-3), "e2e_ms": e2e_ms, "h2d": h2d, "d2h": 16,
            "launches": launches, "clocks": clocks, "losses": [float(l) for l in losses], "n_params": tr.n_params,
            "tflops": flops_line * lines_per_gpu * args.steps * world / (ms * 1e-3) / 1e12, "lines_per_gpu": lines_per_gpu}


def train_line(args, r, world):
    return {"metric": "trained chars/sec (train)", "value": r["value"], "unit": "chars/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r["ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C5 training forward/backward + Adam(clipnorm 5): depth 2 width 512 V 512, %d lines per GPU per step, source = target, "
                                   "80 characters + newline, dropout 0.2, random-init weights (variant U)" % r["lines_per_gpu"],
                       "lines_per_step_per_gpu": r["lines_per_gpu"], "parallelism": "data parallel: NCCL all-reduce (sum) of the flat fp32 gradient vector (%d floats) per step" % r["n_params"],
                       "l2": "inputs and dropout masks differ every step; activations + weights + moments (~0.5 GB) exceed the 126 MB L2"},
            "e2e": {"value": r["e2e_value"], "unit": "chars/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"], "ms_per_step": r["e2e_ms"] / args.steps,
                    "api": "Trainer.train_on_batch (numpy index arrays in, loss out)"},
            "gpu_launches": r["launches"], "clocks": r["clocks"],
            "roofline": {"bound": "latency", "kernel": "the recorded step graph (cuBLAS SGEMM at 64 rows + the hand-written cell / attention / loss kernels)",
                         "achieved": r["tflops"], "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None,
                         "note": "algorithmic 3 x forward FLOPs; at 64 rows per GEMM the step is bound by ~2 500 dependent launches, not by a pipe"},
            "loss_first_last": [r["losses"][0], r["losses"][-1]]}


def make_s2s(topo, w, mapping, device):
    from cor_asv_ann_b200.seq2seq import Sequence2Sequence
    s = Sequence2Sequence()
    s.width, s.depth, s.voc_size, s.mapping = topo["width"], topo["depth"], topo["voc_size"], mapping
    s.device = device
    s.configure()
    for layer, arrs in w.items():
        s.encoder_decoder_model.get_layer(layer).set_weights(arrs)
    s.status = 2
    return s


def measure_c1(device):
    """Config 1: depth 1 / width 128 over the alphabet of the 55 lines of the reference's test pages
    (testdata/.../kant_aufklaerung_1784-complex, frozen in tests/golden/kant_lines.json), one call per page like the
    OCR-D wrapper: plain strings, fast greedy and per-line-greedy semantics."""
    d = json.load(open(os.path.join(ROOT, "tests", "golden", "kant_lines.json"), encoding="utf-8"))
    alphabet, pages = d["alphabet"], d["pages"]
    topo = dict(width=128, depth=1, voc_size=len(alphabet))
    mapping = ({c: i for i, c in enumerate(alphabet)}, {i: c for i, c in enumerate(alphabet)})
    s = make_s2s(topo, build_weights(topo, "U", seed=11), mapping, device)
    plain = [[''.join(chunk[0][0] for chunk in line).replace('\a', '') for line in page] for page in pages]
    n_chars = sum(len(l) - 1 for page in plain for l in page)
    for page in plain:
        s.correct_lines(page, None, fast=True, greedy=True)
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        for page in plain:
            s.correct_lines(page, None, fast=True, greedy=True)
    dt = (time.perf_counter() - t0) / reps
    return {"workload": "C1 depth 1 width 128, the 55 kant test lines, one fast-greedy correct_lines call per page (strings in, strings + probabilities + alignments out)",
            "lines": sum(len(p) for p in plain), "pages": len(plain), "chars": n_chars, "ms_all_pages": dt * 1e3, "chars_per_s": n_chars / dt}


def measure_predict(args, topo, w, mapping, device, n_lines):
    """Row N2, the call a user of the reference makes (scripts/proc.py:62): Sequence2Sequence.predict on a
    text file, fast mode, strings in -> strings + scores out; 64 reference batches per engine call with
    the next super-batch read and encoded in a worker thread, against the reference's one-batch loop."""
    import tempfile
    s = make_s2s(topo, w, mapping, device)
    VOC = topo["voc_size"]
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
    s.rejection_threshold, s.beam_width_in = BEAM_C2["rejection_threshold"], BEAM_C2["beam_width_in"]
    import gc
    import torch
    for fast in (True, False):
        for _ in range(2):
            s.correct_lines(page, None, fast=fast, greedy=fast)
        gc.collect()                    # the predict legs' pinned result buffers are freed here, not inside the timed calls
        torch.cuda.synchronize()
        times = []
        for _ in range(7):
            t0 = time.perf_counter()
            res = s.correct_lines(page, None, fast=fast, greedy=fast)
            times.append((time.perf_counter() - t0) * 1e3)
        assert len(res[0]) == len(page)
        out["page32_correct_lines_ms_%s" % ("fast" if fast else "beam")] = float(np.median(times))
        out["page32_correct_lines_ms_%s_all" % ("fast" if fast else "beam")] = [round(t, 2) for t in times]
    return out


def measure_strong(args, topo, w, mapping, local, rank, world):
    """Strong scaling of the LITERAL configs: C3 = 1 000 000 lines fast greedy, C2 = 100 000 lines beamed, a fixed job
    partitioned over the ranks by whole reference batches (shard.partition_batches), decoded through the product's
    own sharding (shard.decode_job: host strings -> index encoding -> engine -> result arrays), all results on rank 0
    in input order INSIDE the timed region."""
    import torch
    import torch.distributed as dist
    from cor_asv_ann_b200 import shard
    from cor_asv_ann_b200.synthetic import indices_to_lines
    s = make_s2s(topo, w, mapping, local)
    s.beam_budget_bytes = int(args.beam_budget_gb * (1 << 30))
    s.rejection_threshold, s.beam_width_in = BEAM_C2["rejection_threshold"], BEAM_C2["beam_width_in"]
    out = {}
    for name, n_lines, fast, seed in (("C3_greedy_1M_lines", args.strong_greedy_lines, True, 1003), ("C2_beam_100k_lines", args.strong_beam_lines, False, 1002)):
        if n_lines <= 0:
            continue
        # every rank materialises only its own part of the job (as it would read its part of a file)
        parts = shard.partition_batches(n_lines, REF_BATCH, world)
        lo, hi = (parts[rank][0][0], parts[rank][-1][1]) if parts[rank] else (0, 0)

        mine = indices_to_lines(make_indices(hi - lo, seed + 31 * rank), mapping[1]) if hi > lo else []
        job = shard.LocalPart(n_lines, lo, mine)
        runner = shard.engine_runner(s, fast=fast, greedy=fast)
        # warm-up on every rank THROUGH the collectives: NCCL sets up its point-to-point channels on first use (the gather),
        # and the engine records its loop graph for the super-batch geometry
        shard.decode_job(runner, indices_to_lines(make_indices(world * 64 * REF_BATCH, 9), mapping[1]), batch_size=REF_BATCH,
                         superbatch=64, rank=rank, world_size=world)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        res = shard.decode_job(runner, job, batch_size=REF_BATCH, superbatch=64, rank=rank, world_size=world)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if rank == 0:
            assert len(res) == n_lines
            out[name] = {"lines": n_lines, "seconds": float(dt[0]), "chars_per_s": LINE_LEN * n_lines / float(dt[0]),
                         "result": "chars/lens/scores arrays of all lines on rank 0 in input order (strings on demand)",
                         "api": "cor_asv_ann_b200.shard.decode_job(engine_runner(Sequence2Sequence)): strings -> arrays, 64 reference batches per engine call"}
    return out


def bench_line(args, r, workload, world, peaks):
    peak = peaks.get("bf16_tflops_sustained", 1400.0)
    traffic, traffic_src = measured_traffic(workload, r["L"])
    peak_src = "MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks else "fallback 1400 (of fallback)"
    achieved = r["k_flops"] / (r["k_ms"] * 1e-3) / 1e12 if r["k_ms"] else None
    whole = r["algo_flops_line"] * r["L"] * args.steps * world / (r["ms"] * 1e-3) / 1e12
    line = {
        "metric": "corrected chars/sec (%s)" % workload, "value": r["value"], "unit": "chars/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"] / args.steps,
        "ms_each_step": r["ms_each"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16" if args.gemm_mode == 3 else "f32", "data": "synthetic", "config": workload_config(workload, r["L"]),
        "e2e": {"value": r["e2e_value"], "unit": "chars/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                "ms_per_step": r["e2e_ms"] / args.steps, "api": r["api"]},
        "gpu_launches": r["launches"],
        "host_launch_calls": r["host_launches"],     # the decoder loops are device-side WHILE graphs: one graph launch runs all iterations
        "clocks": r["clocks"],
        "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel<1,2,256,LstmEpi>: GEMM + fused LSTM-cell epilogue (CTA pairs)", "achieved": achieved,
                     "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if (peak and achieved) else None, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "launches_per_step": r["k_n"], "kernel_ms_per_step": r["k_ms"],
                     "kernel_share_of_step": r["k_ms"] / r["prof_ms"] if r["prof_ms"] else None,
                     "measured_in": "a separate pass with every launch of that kernel bracketed by CUDA events (not part of `value`)",
                     "whole_step_tflops": whole, "whole_step_frac": whole / peak if peak else None,
                     "precision": {1: "fp32 via FFMA", 3: "fp16 single pass: the reduced-precision VARIANT with the looser tolerance stated in DESIGN.md section 3, not the product default"}.get(args.gemm_mode, "fp32-grade three-pass fp16 hi/lo (ceiling 1/3 of the bf16 peak; see DESIGN.md)")},
        "algorithmic_gflop_per_line": r["algo_flops_line"] / 1e9,
    }
    line.update(r["extra"])
    if r["parity"] is not None:
        line["parity"] = r["parity"]
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=os.environ.get("COR_BENCH_WORKLOAD", "beam"), choices=sorted(WORKLOADS) + ["train"])
    ap.add_argument("--lines", type=int, default=int(os.environ.get("COR_BENCH_LINES", "0")), help="lines per step per GPU (0 = the workload's default)")
    ap.add_argument("--beam-budget-gb", type=float, default=72.0)
    ap.add_argument("--gemm-mode", type=int, default=0)
    ap.add_argument("--cpu-lines", type=int, default=None)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of the last timed step")
    ap.add_argument("--page-lines", type=int, default=32, help="lines of the small-batch (one page) latency leg; 0 = skip")
    ap.add_argument("--page-only", action="store_true", help="run just the small-batch latency leg")
    ap.add_argument("--predict-lines", type=int, default=65536, help="lines of the Sequence2Sequence.predict leg; 0 = skip")
    ap.add_argument("--predict-only", action="store_true", help="run just the predict leg")
    ap.add_argument("--only", action="store_true", help="measure only --workload (skip every other leg)")
    ap.add_argument("--strong-greedy-lines", type=int, default=1000000, help="fixed job of the strong-scaling leg, C3 (0 = skip)")
    ap.add_argument("--strong-beam-lines", type=int, default=100000, help="fixed job of the strong-scaling leg, C2 (0 = skip)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from cor_asv_ann_b200.engine import Engine
    from cor_asv_ann_b200.synthetic import synthetic_alphabet

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
    engines = {}

    def engine_for(workload):
        spec = WORKLOADS[workload]
        key = (json.dumps(spec["topo"], sort_keys=True), spec["variant"])
        for k2, e in engines.items():           # one workspace at a time: a beam workspace takes its whole budget
            if k2 != key:
                e.release_workspace()
        if key not in engines:
            engines[key] = Engine(dict(spec["topo"], window_width=5), build_weights(spec["topo"], spec["variant"]),
                                  device=local, eos_index=1, gemm_mode=args.gemm_mode)
        return engines[key]

    if args.workload == "train":
        r = measure_train(args, local, rank, world)
        if world > 1:
            dist.barrier()
        sys.stdout.flush(); os.dup2(saved_stdout, 1)
        if rank == 0:
            line = train_line(args, r, world)
            if world == 1 and not args.no_cpu:
                v, dt = cpu_oracle_train_rate(args.cpu_lines or 64)
                line["cpu_baseline"] = {"value": v, "unit": "chars/s", "cores": host_threads(), "kind": "port",
                                        "sample": "%d lines of the same workload, one step of the torch float64 autograd oracle (Keras/TF unavailable), %.1f s" % (args.cpu_lines or 64, dt)}
            print(json.dumps(line))
        if world > 1:
            dist.destroy_process_group()
        return
    mapping = synthetic_alphabet(C2["voc_size"])
    w_c2 = build_weights(C2, "U")
    if args.predict_only:
        r = measure_predict(args, C2, w_c2, mapping, local, args.predict_lines)
        sys.stdout.flush(); os.dup2(saved_stdout, 1)
        print(json.dumps({"predict": r}))
        return
    if args.page_only:
        print_page = measure_page(args, engine_for("beam"), args.page_lines)
        sys.stdout.flush(); os.dup2(saved_stdout, 1)
        print(json.dumps({"page": print_page}))
        return
    single = world == 1 and not args.only
    par = 0 if (args.no_parity or world > 1) else 1
    lines_of = lambda wl: args.lines or WORKLOADS[wl]["lines"]
    main_r = measure(args, engine_for(args.workload), args.workload, lines_of(args.workload), rank, world, local, sample_clocks=(rank == 0),
                     parity_lines=par * {"beam": 64, "greedy": 256, "beam_competitive": 8, "beam_stress": 1, "c4": 0}[args.workload])
    others = {}
    if not args.only:
        first = "greedy" if args.workload == "beam" else "beam"
        others[first] = measure(args, engine_for(first), first, lines_of(first), rank, world, local, sample_clocks=(rank == 0),
                                parity_lines=par * (256 if first == "greedy" else 64))
    if single:
        short = argparse.Namespace(**dict(vars(args), steps=2, warmup=3))
        for wl, n_par in (("beam_competitive", 8), ("beam_stress", 1), ("c4", 0)):
            if wl != args.workload:
                others[wl] = (short, measure(short, engine_for(wl), wl, WORKLOADS[wl]["lines"], rank, world, local, sample_clocks=False,
                                             roofline_pass=False, parity_lines=par * n_par))
    page_r = measure_page(args, engine_for("beam"), args.page_lines) if (rank == 0 and args.page_lines > 0 and single) else None
    for e in engines.values():                  # the legs below build their own Sequence2Sequence objects (own engines, own workspaces)
        e.release_workspace()
    predict_r = measure_predict(args, C2, w_c2, mapping, local, args.predict_lines) if (rank == 0 and single and args.predict_lines > 0) else None
    c1_r = measure_c1(local) if (rank == 0 and single) else None
    early_r = measure_early_exit(args, local) if (rank == 0 and single) else None
    train_r = None
    if rank == 0 and single:
        for e in engines.values():
            e.release_workspace()
        train_r = train_line(args, measure_train(args, local, 0, 1), 1)
    strong_r = measure_strong(args, C2, w_c2, mapping, local, rank, world) if not args.only else None

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
        keys = ("metric", "value", "unit", "steps", "ms_per_step", "ms_each_step", "config", "e2e", "gpu_launches", "host_launch_calls", "clocks", "roofline",
                "algorithmic_gflop_per_line", "parity", "beam_iterations_per_line", "hypothesis_steps_per_line",
                "hypotheses_per_line_and_iteration", "hypothesis_steps_per_s", "finals_per_line", "fringe_peak_mean", "fringe_peak_max",
                "near_tie_decisions_per_line", "lines_without_final", "lines_equal_to_input", "near_tie_argmaxes_last_step")
        also = {}
        for wl, r in others.items():
            a2, r2 = (r if isinstance(r, tuple) else (args, r))
            o = bench_line(a2, r2, wl, world, peaks)
            also[wl] = {k: o[k] for k in keys if k in o}
        if also:
            line["also"] = also
        if "beam_competitive" in also:
            # the literal config is a one-hypothesis search with random weights (see the module docstring); the search
            # under competition is measured beside it and is the number to read for "what does the beam search cost"
            c = also["beam_competitive"]
            line["co_primary"] = {"workload": "beam_competitive", "value": c["value"], "unit": "chars/s", "e2e": c["e2e"]["value"],
                                  "hypothesis_steps_per_s": c.get("hypothesis_steps_per_s"),
                                  "hypotheses_per_line_and_iteration": c.get("hypotheses_per_line_and_iteration"),
                                  "fringe_peak_max": c.get("fringe_peak_max"), "parity": c.get("parity")}
        if c1_r is not None:
            line.setdefault("also", {})["c1"] = c1_r
        if early_r is not None:
            line.setdefault("also", {})["greedy_early_exit"] = early_r
        if train_r is not None:
            line.setdefault("also", {})["train_c5"] = {k: train_r[k] for k in ("metric", "value", "unit", "ms_per_step", "config", "e2e", "gpu_launches", "roofline", "loss_first_last")}
        if page_r is not None:
            line["page"] = page_r
        if predict_r is not None:
            line["predict"] = predict_r
        if strong_r:
            line["strong"] = dict(strong_r, scaling="strong", n_gpus=world)
        if world == 1 and not args.no_cpu:
            cores = host_threads()
            from threadpoolctl import threadpool_limits
            threadpool_limits(limits=cores)
            n_cpu = args.cpu_lines or {"greedy": 256, "beam": 384}.get(args.workload, 4)   # ~10-20 s of CPU work
            v, dt = cpu_oracle_rate(args.workload, n_cpu)
            line["cpu_baseline"] = {"value": v, "unit": "chars/s", "cores": cores, "kind": "port",
                                    "sample": "%d lines of the same workload, numpy oracle (Keras/TF unavailable), %.1f s"
                                              % (n_cpu, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

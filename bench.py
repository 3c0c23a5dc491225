#!/usr/bin/env python3
"""bench.py -- throughput of the hot path on B200 (contract: see the task statement).

Default workload ("main") = BASELINE.json configs[4], the largest single-GPU configuration and the one decode
shape that is HBM-bound rather than launch-bound: RWKV v4 Pile-14B shape, Q4_1 body + FP16 head, random-init
weights in the ggmf format, single-token decode of one stream with logits.  One "step" = one token through every
layer + the head.
  N = 1   the whole model on one GPU
  N > 1   the same model tensor-parallel over the N ranks (column / row split weights, all-reduce over NVLink
          peer memory twice per layer, logits all-gather): same metric string, "scaling": "strong"
  value         tokens/s with state + token ids resident in HBM (CUDA-graph replay, device-timed; 9.6 GB of weights
                per step, 75 x the L2)
  e2e           tokens/s through the rwkv.h C ABI (rwkv_eval with HOST state / logits buffers, copies inside the
                timed region); pinned buffers, and pageable numpy buffers as the reference's callers pass them
  roofline      the STEP: algorithmic bytes of one token (every weight byte except the embedding rows not gathered,
                + state + vectors) / device time per step, against the measured HBM copy bandwidth
  cpu_baseline  the unmodified reference CPU library (oracle/_ref) on the same model file and tokens; its logits
                are compared with the GPU arm's ("parity_max_rel")
  extras        (N = 1) the other BASELINE configs, one compact record each: 169M decode (configs[1]), the Q5_1
                mat-vec table, v6-1.6B prefill (configs[2]), 256 v7-2.9B streams (configs[3])
  --impl reference   times that reference alone (all host threads), same config dict

Other workloads (`--workload`) run one config alone: decode (169M), matvec, prefill, streams, tp.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PF = ctypes.POINTER(ctypes.c_float)
PU = ctypes.POINTER(ctypes.c_uint32)



def ncu_traffic(key):
    """DRAM bytes per launch of the roofline kernel, from the committed ncu --set full capture
    (profiles/traffic.json); None when no capture is recorded for it."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            return json.load(f)[key]["traffic_bytes"]
    except Exception:
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0))), "measured"
    return 6650.0, 1400.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def reference_lib():
    """The unmodified reference CPU library (oracle/_ref) -- measured as a baseline, never shipped."""
    from oracle.refload import load_reference
    return load_reference()


def time_reference_decode(model_file: str, tokens, warmup: int, per_repeat: int, repeats: int = 5, threads=None, want_logits_at=None):
    """tokens/s of the reference's rwkv_eval loop on the host cores: ONE model load, `warmup` evals, then `repeats`
    timed runs of `per_repeat` evals each (>= 0.5 s per run is the caller's job); the best run counts (ggml spawns a
    disposable thread pool per call, so short cold samples under-report).  Returns (tok_s, threads, sysinfo, logits
    after token index `want_logits_at` or None, all runs)."""
    L = reference_lib()
    L.lib.rwkv_set_print_errors(None, False)
    n = threads or host_cores()
    t_load = time.perf_counter()
    ctx = L.load(model_file, n, 0)
    time_reference_decode.last_load_s = time.perf_counter() - t_load
    st = None
    lg = np.empty(ctx.logits_len, np.float32)
    keep = None
    for i, t in enumerate(tokens[:warmup]):
        st, _ = ctx.eval(int(t), st, state_out=st, logits_out=lg)
        if want_logits_at is not None and i == want_logits_at:
            keep = lg.copy()
    runs = []
    pos = warmup
    for _ in range(repeats):
        seg = [int(t) for t in tokens[pos:pos + per_repeat]]
        if len(seg) < per_repeat:
            pos = warmup
            seg = [int(t) for t in tokens[pos:pos + per_repeat]]
        pos += per_repeat
        t0 = time.perf_counter()
        for t in seg:
            st, _ = ctx.eval(t, st, state_out=st, logits_out=lg)
        runs.append(len(seg) / (time.perf_counter() - t0))
    ctx.free()
    return max(runs), n, L.system_info(), keep, runs


class Dist:
    """torch.distributed only as plumbing: barrier + max over ranks.  One process group per process: Dist() returns the same object
    every time (the main workload and its extras share it); close() only takes effect through Dist.shutdown()."""
    _inst = None

    def __new__(cls, use_cuda: bool = True):
        if cls._inst is None:
            cls._inst = super().__new__(cls)
            cls._inst._setup(use_cuda)
        return cls._inst

    def __init__(self, use_cuda: bool = True):
        pass

    def _setup(self, use_cuda: bool = True):
        self.rank, self.world, self.local = dist_env()
        self.torch = None
        if self.world > 1 or use_cuda:
            import torch
            self.torch = torch
            if use_cuda and torch.cuda.is_available():
                torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            backend = "nccl" if (use_cuda and self.torch.cuda.is_available()) else "gloo"
            dist.init_process_group(backend=backend, rank=self.rank, world_size=self.world)
            self.dist = dist
            self.device = self.torch.device("cuda", self.local) if backend == "nccl" else self.torch.device("cpu")

    def barrier(self):
        if self.torch is not None and self.torch.cuda.is_available():
            self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        if self.torch is not None and self.torch.cuda.is_available():
            self.torch.cuda.synchronize()

    def max(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        pass

    @classmethod
    def shutdown(cls):
        if cls._inst is not None and cls._inst.world > 1:
            cls._inst.dist.destroy_process_group()
        cls._inst = None


def head_roofline(lib, n_vocab: int, n_embed: int, iters: int = 200):
    """The dominant kernel of a 169M decode step: the FP16 head mat-vec (V x D x 2 bytes).
    Timed with CUDA events on the engine's stream, rotating over 4 device copies (308 MB >
    the 126 MB L2), so every launch streams its matrix from HBM."""
    rng = np.random.default_rng(0)
    w = (rng.standard_normal((n_vocab, n_embed)) / np.sqrt(n_embed)).astype(np.float16)
    m = lib.library.rwkv_b200_matrix_create(1, w.ctypes.data, n_vocab, n_embed, 4)
    if not m:
        raise RuntimeError("rwkv_b200_matrix_create failed")
    ms = ctypes.c_double(0.0)
    ok = lib.library.rwkv_b200_matrix_bench(m, 1, iters, ctypes.byref(ms))
    nbytes = lib.library.rwkv_b200_matrix_bytes(m) + n_embed * 4 + n_vocab * 4
    lib.library.rwkv_b200_matrix_free(m)
    if not ok:
        raise RuntimeError("rwkv_b200_matrix_bench failed")
    return nbytes, ms.value


DECODE_CFG = {"workload": "RWKV v4 Pile-169M shape (L12 D768 V50277 F3072), Q4_0 body + FP16 head, single-token decode with logits, "
                          "1 stream per GPU (BASELINE.json configs[1])",
              "shape": "v4-169m", "format": "FP16-to-Q4_0", "streams_per_gpu": 1}


def run_decode(args):
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    d = Dist()
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    if lib.library.rwkv_b200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU path")
    lib.library.rwkv_b200_set_device(d.local)
    if d.rank == 0:
        path = synth.ensure_model(lib, DECODE_CFG["shape"], DECODE_CFG["format"])
    d.barrier()
    path = synth.ensure_model(lib, DECODE_CFG["shape"], DECODE_CFG["format"])
    model = rwkv_cppy_b200.RWKVModel(lib, path)
    V, S = model.n_vocab, model.state_len
    K, W = args.steps, max(args.warmup, 3)
    rng = np.random.default_rng(1 + d.rank)
    toks = rng.integers(0, V, size=(W + K, 1)).astype(np.uint32)

    # ---- value: state + tokens resident, CUDA-graph replay, device-timed, cold L2 per step ----
    batch = model.create_batch(1)
    launches = batch.launches_per_step(True)
    batch.run_timed(toks[:W], True)
    d.barrier()
    ms_warm = batch.run_timed(toks[W:], True)                 # L2 as the decode loop leaves it
    lib.library.rwkv_b200_set_l2_flush(256 << 20)
    batch.run_timed(toks[:W], True)
    d.barrier()
    sampler = ClockSampler(d.local)
    sampler.start()
    ms = batch.run_timed(toks[W:], True)                      # every step starts with a flushed L2
    d.barrier()
    clocks = sampler.stop()
    lib.library.rwkv_b200_set_l2_flush(0)
    ms_max = d.max(ms)
    ms_warm_max = d.max(ms_warm)
    batch.free()

    # ---- e2e: the rwkv.h call a user makes, host buffers, copies inside the timed region ----
    import torch
    state_t = torch.zeros(S, dtype=torch.float32).pin_memory()   # pinned host buffers (bench contract); the ABI takes any host pointer
    logits_t = torch.zeros(V, dtype=torch.float32).pin_memory()
    state, logits = state_t.numpy(), logits_t.numpy()
    L = lib.library
    ctxp = model._ctx.ptr
    L.rwkv_init_state(ctxp, state.ctypes.data_as(PF))
    sp, lp = state.ctypes.data_as(PF), logits.ctypes.data_as(PF)
    for t in toks[:W, 0]:
        assert L.rwkv_eval(ctxp, int(t), sp, sp, lp)
    d.barrier()
    t0 = time.perf_counter()
    for t in toks[W:, 0]:
        L.rwkv_eval(ctxp, int(t), sp, sp, lp)
    e2e_s = time.perf_counter() - t0
    d.barrier()
    e2e_s = d.max(e2e_s)
    assert np.isfinite(logits).all()

    out = None
    if d.rank == 0:
        hbm, _, which = peaks()
        nbytes, kms = head_roofline(lib, V, model.n_embed)
        ach = nbytes / (kms * 1e-3) / 1e9
        out = {
            "metric": "decode tokens/s (RWKV v4 169M-shape Q4_0, single stream per GPU)", "value": d.world * K / (ms_max * 1e-3),
            "unit": "tokens/s", "n_gpus": d.world, "steps": K, "warmup": W, "ms_per_step": ms_max / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 dp4a blocks + f32 accumulate (Q4_0 body), f16 head",
            "data": "synthetic (random-init weights in ggmf format, uniform random token ids)",
            "config": DECODE_CFG,
            "detail": {"l2": "256 MiB scratch written before every timed step (cold L2), one CUDA event pair per step",
                       "value_l2_warm": d.world * K / (ms_warm_max * 1e-3)},
            "clocks": clocks,
            "e2e": {"value": d.world * K / e2e_s, "unit": "tokens/s", "h2d_bytes_per_step": S * 4 + 4, "d2h_bytes_per_step": S * 4 + V * 4,
                    "api": "rwkv_eval(ctx, token, state_in, state_out, logits_out) with pinned host buffers"},
            "gpu_launches": launches * K,
            # the STEP against the HBM roofline (129.8 MB per token); this shape is launch-latency-bound, not HBM-bound
            "roofline": {"bound": "hbm", "achieved": 129.8e6 / (ms_max / K * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": 129.8e6 / (ms_max / K * 1e-3) / 1e9 / hbm, "traffic": None,
                         "what": "whole decode step: 129.8 MB algorithmic bytes per token / device time per token",
                         "largest_kernel": {"kernel": "matvec_kernel<F16,1> head 50277x768 (77.2 MB)", "us_per_launch": kms * 1e3, "achieved": ach, "frac": ach / hbm,
                                            "traffic": ncu_traffic("head_gemv_f16_50277x768")},
                         "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs)" if which == "measured" else which},
        }
        if not args.no_cpu:
            try:
                tok_s, thr, info, _, runs = time_reference_decode(path, toks[:, 0], 8, 64, repeats=5, threads=min(host_cores(), 16))
                out["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference",
                                       "sample": "best of 5 runs of 64 rwkv_eval calls (after 8 warm-up, one model load) of the unmodified reference CPU build on the same model file; " + info}
            except Exception as e:  # the baseline is a report, not a dependency
                out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": 0, "kind": "reference", "sample": f"not run: {e}"}
    model.free()
    d.close()
    return out


MAIN_SHAPE, MAIN_FMT = "v4-14b", "Q4_1"
MAIN_CFG = {"workload": "RWKV v4 Pile-14B shape (L40 D5120 V50277 F20480), Q4_1 body + FP16 head, single-token decode of one stream with logits "
                        "(BASELINE.json configs[4]); one GPU holds the whole model, N > 1 GPUs hold it tensor-parallel",
            "shape": MAIN_SHAPE, "format": MAIN_FMT, "streams": 1}
MAIN_METRIC = "decode tokens/s (RWKV v4 14B-shape Q4_1, 1 stream)"
REF_WARMUP, REF_PER_REPEAT, REF_REPEATS = 2, 5, 5          # reference CPU arm: ~0.15 s per token on 14B -> >= 0.5 s per timed run


def main_tokens(n_vocab: int, n: int):
    return np.random.default_rng(1).integers(0, n_vocab, size=n).astype(np.uint32)


def reference_leg(path: str, n_vocab: int):
    """The reference CPU implementation on the main workload: shared by --impl reference and the cpu_baseline leg, so the two
    agree by construction (same threads, tokens, repeats)."""
    toks = main_tokens(n_vocab, REF_WARMUP + REF_PER_REPEAT * REF_REPEATS)
    tok_s, thr, info, lg, runs = time_reference_decode(path, toks, REF_WARMUP, REF_PER_REPEAT, REF_REPEATS, want_logits_at=REF_WARMUP - 1)
    sample = (f"best of {REF_REPEATS} runs of {REF_PER_REPEAT} rwkv_eval calls (one model load, {REF_WARMUP} warm-up evals) of the unmodified reference "
              f"(oracle/_ref) with {thr} threads on the same model file and token stream; runs tok/s = {[round(r, 2) for r in runs]}; "
              f"rwkv_init_from_file took {time_reference_decode.last_load_s:.2f} s; " + info)
    return tok_s, thr, sample, lg


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the main workload (rank 0 only)."""
    rank, world, _ = dist_env()
    if rank != 0:
        return None
    from rwkv_cppy_b200 import synth                           # pure-numpy writer: this process never maps librwkv.so
    path = synth.ensure_model(None, MAIN_SHAPE, MAIN_FMT)
    tok_s, thr, sample, _ = reference_leg(path, synth.SHAPES[MAIN_SHAPE].n_vocab)
    return {"impl": "reference", "metric": MAIN_METRIC, "value": tok_s, "unit": "tokens/s",
            "n_gpus": world, "steps": REF_PER_REPEAT, "warmup": REF_WARMUP, "ms_per_step": 1e3 / tok_s, "higher_is_better": True,
            "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "int8 blocks + f32 accumulate (ggml CPU)", "data": "synthetic (random-init weights in ggmf format, uniform random token ids)",
            "config": MAIN_CFG,
            "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference", "sample": sample},
            "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main_bytes_per_gpu(shape, world: int) -> dict:
    """Algorithmic bytes one decode step moves per GPU: every quantised block and FP16 head row this rank owns
    once, the replicated FP32 vectors, one embedding row, the v4 state read + written."""
    D, L, V, F = shape.n_embed, shape.n_layer, shape.n_vocab, shape.ffn
    body = (4 * D * D + 2 * D * F + D * D) * L * 20 // 32          # att k, v, r, out + ffn key, value, receptance; Q4_1 = 20 bytes per 32 weights
    head = V * D * 2
    vectors = (11 * D * L + 4 * D) * 4                               # ln1, ln2 (w, b), 5 mixes, time_first, time_decay per layer; ln0, ln_out
    state = 2 * 5 * D * L * 4
    emb_row = D * 2
    return {"body": body // world, "head": head // world, "vectors": vectors, "state": state, "emb_row": emb_row,
            "total": body // world + head // world + vectors + state + emb_row}


def kernel_split():
    """Share of a decode step by kernel class, from the committed ncu launch list of this workload (profiles/)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "main_kernel_split.json")))
    except Exception:
        return None


def run_main(args):
    """BASELINE configs[4]: v4 14B-shape Q4_1 single-stream decode; one GPU, or tensor parallel over the ranks."""
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    d = Dist()
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    if lib.library.rwkv_b200_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU path")
    lib.library.rwkv_b200_set_device(d.local)
    t0 = time.perf_counter()
    if d.rank == 0:
        synth.ensure_model(None, MAIN_SHAPE, MAIN_FMT)
    d.barrier()
    path = synth.ensure_model(None, MAIN_SHAPE, MAIN_FMT)
    synth_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    if d.world > 1:
        ids = [lib.rwkv_b200_tp_unique_id() if d.rank == 0 else None]
        d.dist.broadcast_object_list(ids, src=0)
        model = rwkv_cppy_b200.RWKVModel(lib, path, tensor_parallel=(d.rank, d.world, ids[0]))
        if os.environ.get("RWKV_B200_TP_NCCL_ONLY") != "1":      # collectives over NVLink peer memory instead of NCCL
            handles = [None] * d.world
            d.dist.all_gather_object(handles, lib.rwkv_b200_tp_ipc_handle(model._ctx))
            lib.rwkv_b200_tp_connect(model._ctx, handles)
    else:
        model = rwkv_cppy_b200.RWKVModel(lib, path)
    load_s = time.perf_counter() - t0
    V, S = model.n_vocab, model.state_len
    K, W = args.steps, max(args.warmup, 3)
    toks = main_tokens(V, W + K).reshape(-1, 1)                  # the same stream on every rank

    # ---- value: state + tokens resident, CUDA-graph replay, device-timed ----
    batch = model.create_batch(1)
    launches = batch.launches_per_step(True)
    batch.run_timed(toks[:W], True)
    d.barrier()
    sampler = ClockSampler(d.local)
    sampler.start()
    ms = batch.run_timed(toks[W:], True)
    d.barrier()
    clocks = sampler.stop()
    ms_max = d.max(ms)
    batch.free()

    # ---- e2e: rwkv_eval through the C ABI, host buffers, copies inside the timed region (every rank calls it: collective under TP) ----
    import torch
    L = lib.library
    ctxp = model._ctx.ptr
    n_e2e = min(K, 32)

    def e2e_run(state, logits):
        L.rwkv_init_state(ctxp, state.ctypes.data_as(PF))
        sp, lp = state.ctypes.data_as(PF), logits.ctypes.data_as(PF)
        for t in toks[:W, 0]:
            assert L.rwkv_eval(ctxp, int(t), sp, sp, lp)
        d.barrier()
        t0 = time.perf_counter()
        for t in toks[W:W + n_e2e, 0]:
            L.rwkv_eval(ctxp, int(t), sp, sp, lp)
        dt = time.perf_counter() - t0
        d.barrier()
        assert np.isfinite(logits).all()
        return d.max(dt)

    e2e_pinned = e2e_run(torch.zeros(S, dtype=torch.float32).pin_memory().numpy(), torch.zeros(V, dtype=torch.float32).pin_memory().numpy())
    e2e_pageable = e2e_run(np.zeros(S, np.float32), np.zeros(V, np.float32))          # what the reference's callers pass (rwkv_cpp_model.py:330-351)

    # logits after the reference leg's comparison token, for the parity figure
    gpu_lg = None
    if d.world == 1:
        ref_toks = main_tokens(V, REF_WARMUP + REF_PER_REPEAT * REF_REPEATS)
        st = np.zeros(S, np.float32); gpu_lg = np.zeros(V, np.float32)
        L.rwkv_init_state(ctxp, st.ctypes.data_as(PF))
        for t in ref_toks[:REF_WARMUP]:
            L.rwkv_eval(ctxp, int(t), st.ctypes.data_as(PF), st.ctypes.data_as(PF), gpu_lg.ctypes.data_as(PF))
    model.free()

    out = None
    if d.rank == 0:
        hbm, _, which = peaks()
        step_ms = ms_max / K
        nb = main_bytes_per_gpu(synth.SHAPES[MAIN_SHAPE], d.world)
        ach = nb["total"] / (step_ms * 1e-3) / 1e9
        out = {
            "metric": MAIN_METRIC, "value": K / (ms_max * 1e-3), "unit": "tokens/s", "n_gpus": d.world, "steps": K, "warmup": W,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong" if d.world > 1 else "weak", "vs_baseline": None,
            "dtype": "int8 dp4a blocks + f32 accumulate (Q4_1 body), f16 head", "data": "synthetic (random-init weights in ggmf format, uniform random token ids)",
            "config": MAIN_CFG, "clocks": clocks,
            "e2e": {"value": n_e2e / e2e_pinned, "unit": "tokens/s", "h2d_bytes_per_step": S * 4 + 4, "d2h_bytes_per_step": S * 4 + V * 4,
                    "api": "rwkv_eval(ctx, token, state_in, state_out, logits_out), pinned host buffers" + (" on every rank" if d.world > 1 else ""),
                    "value_pageable_buffers": n_e2e / e2e_pageable},
            "gpu_launches": launches * K,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": ncu_traffic("main_step"),
                         "what": "the whole decode step on one rank: algorithmic bytes per GPU per token / device time per token "
                                 "(weights stream once per token; 75 x the L2, so every step is cold)",
                         "algorithmic_bytes_per_gpu_per_step": nb, "kernel_split": kernel_split(),
                         "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs)" if which == "measured" else which},
            "detail": {"parallelism": f"tp{d.world}" if d.world > 1 else "1 gpu", "launches_per_step": launches, "model_synth_s": round(synth_s, 1),
                       "model_load_s": round(load_s, 2), "file_bytes": os.path.getsize(path),
                       "collective": None if d.world == 1 else ("NCCL" if os.environ.get("RWKV_B200_TP_NCCL_ONLY") == "1" else
                                     ("fused: push from the mat-vec epilogue + gather in the next layer norm" if os.environ.get("RWKV_B200_TP_FUSE") == "1"
                                      else "low-latency all-reduce kernel over NVLink peer memory: {value, call counter} entries pushed into every peer's inbox, summed in rank order (csrc/tp_p2p.cu)"))},
        }
        if d.world == 1 and not args.no_cpu:
            try:
                tok_s, thr, sample, ref_lg = reference_leg(path, V)
                out["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference", "sample": sample}
                if ref_lg is not None and gpu_lg is not None:
                    out["cpu_baseline"]["parity_max_rel"] = float(np.abs(gpu_lg - ref_lg).max() / max(1.0, float(np.abs(ref_lg).max())))
                    out["cpu_baseline"]["parity_note"] = f"max |logit_gpu - logit_reference| / max |logit_reference| after {REF_WARMUP} tokens of the same stream"
            except Exception as e:  # the baseline is a report, not a dependency
                out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": 0, "kind": "reference", "sample": f"not run: {e}"}
    if d.world == 1 and out is not None and not args.no_extras:
        out["extras"] = run_extras(args)
    elif d.world > 1 and not args.no_extras:
        # BASELINE configs[3] on the same ranks: 256 independent v7 2.9B-shape streams sharded 256 / N per GPU, weights replicated, no
        # communication in the data path.  Every rank runs it; rank 0 reports.
        ns = argparse.Namespace(**vars(args))
        ns.steps, ns.warmup, ns.streams, ns.shape, ns.no_cpu = 16, 3, 256, None, True
        try:
            r = run_streams(ns)
            rec = None if r is None else {"config": f"configs[3]: v7 2.9B-shape Q8_0, 256 independent streams sharded over {d.world} GPUs ({256 // d.world} per GPU), no collective",
                                          "tokens_per_s": r["value"], "ms_per_step": r["ms_per_step"], "scaling": "strong (fixed 256-stream job)",
                                          "step_roofline_frac_per_gpu": r["roofline"]["frac"], "step_hbm_gbs_per_gpu": r["roofline"]["achieved"]}
        except Exception as e:
            rec = {"error": str(e)[:300]}
        if out is not None:
            out["extras"] = {"streams_v7_2b9_sharded": rec}
    return out


def run_extras(args):
    """The other BASELINE configs at N = 1, one compact record each (each is also a workload of its own: --workload decode | matvec | prefill | streams)."""
    ex = {}
    ns = argparse.Namespace(**vars(args))
    ns.no_cpu = True

    def attempt(name, fn):
        t0 = time.perf_counter()
        try:
            ex[name] = fn()
        except Exception as e:                                    # an extra never fails the headline line
            ex[name] = {"error": str(e)[:300]}
        ex[name]["wall_s"] = round(time.perf_counter() - t0, 1)

    def decode():
        ns.steps, ns.warmup = 256, 16
        r = run_decode(ns)
        return {"config": "configs[1]: v4 169M-shape Q4_0 decode, 1 stream", "tokens_per_s": r["value"], "us_per_step": r["ms_per_step"] * 1e3,
                "e2e_tokens_per_s": r["e2e"]["value"], "launches_per_step": r["gpu_launches"] // r["steps"],
                "step_roofline_frac": r["roofline"]["frac"], "step_hbm_gbs": r["roofline"]["achieved"]}

    def matvec():
        r = run_gemv(ns)
        return {"config": "headline kernel metric: dequant-fused mat-vec, rotating copies > L2", "peak_gbs": r["roofline"]["peak"],
                "rows": [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items() if k in ("type", "M", "K", "us", "GBps", "frac_measured")}
                         for row in r["config"]["rows"]]}

    def prefill():
        ns.steps, ns.warmup, ns.tokens, ns.shape = 5, 2, 1024, None
        r = run_prefill(ns)
        return {"config": "configs[2]: v6 1.6B-shape Q5_1 rwkv_eval_sequence T=1024", "tokens_per_s": r["value"], "ms_per_pass": r["ms_per_step"],
                "e2e_tokens_per_s": r["e2e"]["value"], "whole_pass_tflops": r["roofline"]["whole_pass_tflops"],
                "whole_pass_tensor_frac": r["roofline"]["whole_pass_tflops"] / r["roofline"]["peak"],
                "ffn_key_matmul_tflops": r["roofline"]["achieved"], "ffn_key_matmul_frac": r["roofline"]["frac"]}

    def streams():
        ns.steps, ns.warmup, ns.streams, ns.shape = 16, 3, 256, None
        r = run_streams(ns)
        return {"config": "configs[3]: v7 2.9B-shape Q8_0, 256 independent streams on one GPU", "tokens_per_s": r["value"], "ms_per_step": r["ms_per_step"],
                "e2e_tokens_per_s": r["e2e"]["value"], "step_roofline_frac": r["roofline"]["frac"], "step_hbm_gbs": r["roofline"]["achieved"]}

    attempt("decode_169m", decode)
    attempt("matvec", matvec)
    attempt("prefill_v6_1b6", prefill)
    attempt("streams_v7_2b9", streams)
    return ex


GEMV_SHAPES = [("Q5_1", 8, 5120, 5120), ("Q5_1", 8, 20480, 5120), ("Q5_1", 8, 5120, 20480), ("Q5_1", 8, 7168, 2048),
               ("Q4_0", 2, 20480, 5120), ("Q8_0", 9, 20480, 5120), ("FP16", 1, 65536, 2048), ("FP16", 1, 50277, 768)]


def run_gemv(args):
    """BASELINE headline kernel metric: achieved HBM GB/s of the dequant-fused mat-vec."""
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    hbm, _, which = peaks()
    rng = np.random.default_rng(0)
    rows = []
    shapes = GEMV_SHAPES if args.only is None else [GEMV_SHAPES[args.only]]
    for name, tid, M, K in shapes:
        rows_gen = min(M, 4096)
        if tid == 1:
            raw1 = ((rng.random((rows_gen, K), dtype=np.float32) * 2 - 1) / np.float32(np.sqrt(K))).astype(np.float16).view(np.uint8).reshape(-1)
        else:
            raw1 = synth._random_blocks(rng, name, rows_gen * K // 32, float(1 / np.sqrt(K))).reshape(-1)      # valid ggml blocks of the format
        raw = np.ascontiguousarray(np.tile(raw1, (M + rows_gen - 1) // rows_gen)[: (raw1.size // rows_gen) * M])
        one = raw.size
        copies = max(2, int(np.ceil(400e6 / one)))
        m = lib.library.rwkv_b200_matrix_create(tid, raw.ctypes.data, M, K, copies)
        ms = ctypes.c_double(0.0)
        assert m and lib.library.rwkv_b200_matrix_bench(m, 1, 200, ctypes.byref(ms))
        nbytes = one + K * 4 + M * 4
        lib.library.rwkv_b200_matrix_free(m)
        gbs = nbytes / (ms.value * 1e-3) / 1e9
        rows.append({"type": name, "M": M, "K": K, "MB": one / 1e6, "us": ms.value * 1e3, "GBps": gbs, "frac_measured": gbs / hbm, "frac_8TBps": gbs / 8000.0, "copies": copies})
        print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
    best = max(r["GBps"] for r in rows if r["type"] == "Q5_1")
    return {"metric": "Q5_1 mat-vec achieved HBM GB/s (best of the 14B / 1.6B shapes)", "value": best, "unit": "GB/s", "n_gpus": 1, "steps": 200, "warmup": 4,
            "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 dp4a blocks + f32 accumulate",
            "data": "synthetic", "config": {"workload": "mat-vec microbenchmark, rotating copies > L2", "rows": rows, "peak_source": which},
            "roofline": {"bound": "hbm", "achieved": best, "peak": hbm, "unit": "GB/s", "frac": best / hbm, "traffic": ncu_traffic("gemv_async_q5_1_20480x5120")}}


def run_prefill(args):
    """BASELINE configs[2]: v6 1.6B-shape Q5_1, rwkv_eval_sequence of 1024 tokens."""
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    d = Dist()
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    lib.library.rwkv_b200_set_device(d.local)
    shape = args.shape or "v6-1b6"
    if d.rank == 0:
        synth.ensure_model(lib, shape, "Q5_1")
    d.barrier()
    path = synth.ensure_model(lib, shape, "Q5_1")
    model = rwkv_cppy_b200.RWKVModel(lib, path)
    T = args.tokens
    toks = np.random.default_rng(1 + d.rank).integers(0, model.n_vocab, size=T).astype(np.uint32)
    ms = ctypes.c_double(0.0)
    tp = toks.ctypes.data_as(PU)
    assert lib.library.rwkv_b200_sequence_run_timed(model._ctx.ptr, tp, T, 1, max(args.warmup, 1), ctypes.byref(ms))
    d.barrier()
    sampler = ClockSampler(d.local); sampler.start()
    assert lib.library.rwkv_b200_sequence_run_timed(model._ctx.ptr, tp, T, 1, args.steps, ctypes.byref(ms))
    d.barrier()
    clocks = sampler.stop()
    ms_max = d.max(ms.value)
    state = np.zeros(model.state_len, np.float32); logits = np.zeros(model.n_vocab, np.float32)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        assert lib.library.rwkv_eval_sequence(model._ctx.ptr, toks.view(np.int32).ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), T, None,
                                              state.ctypes.data_as(PF), logits.ctypes.data_as(PF))
    e2e = (time.perf_counter() - t0) / args.steps
    out = None
    if d.rank == 0:
        # algorithmic FLOPs of one pass: 2*T*(every matrix element the sequence multiplies) + the head for the last token
        sh = synth.SHAPES[shape]
        D, F, L, V = sh.n_embed, sh.ffn, sh.n_layer, sh.n_vocab
        per_tok = (5 * D * D + 2 * D * F + D * D + 2 * 5 * sh.mix_lora * D + 2 * sh.decay_lora * D) if sh.arch == 6 else (4 * D * D + 2 * D * F + (D * D if sh.arch != 7 else 0))
        flops = 2.0 * T * per_tok * L + 2.0 * D * V
        _, tf_peak, which = peaks()
        # dominant kernel: the ffn-key GEMM (F x D weights, T rows), timed live on the launching stream with CUDA events
        roof = {"bound": "tensor", "achieved": None, "peak": tf_peak, "unit": "TFLOP/s", "frac": None, "traffic": None,
                "peak_source": which + " (MEASURED_PEAKS.json dense bf16, sustained)", "whole_pass_tflops": flops / (ms_max * 1e-3) / 1e12}
        try:
            raw1 = synth._random_blocks(np.random.default_rng(0), "Q5_1", 256 * D // 32, float(1 / np.sqrt(D))).reshape(-1)
            raw = np.ascontiguousarray(np.tile(raw1, (F + 255) // 256)[: (raw1.size // 256) * F])
            L_ = lib.library
            L_.rwkv_b200_matrix_create.restype = ctypes.c_void_p
            L_.rwkv_b200_matrix_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t]
            mh = L_.rwkv_b200_matrix_create(8, raw.ctypes.data, F, D, 4)
            msl = ctypes.c_double(0)
            if mh and L_.rwkv_b200_matrix_bench_tc(ctypes.c_void_p(mh), T, 50, ctypes.byref(msl)):
                ach = 2.0 * F * D * T / (msl.value * 1e-3) / 1e12
                roof.update({"achieved": ach, "frac": ach / tf_peak, "us_per_launch": msl.value * 1e3,
                             "kernel": f"gemm_tc_kernel<Q5_1> ffn key {F}x{D} x {T} rows (tcgen05.mma kind::f16, FP32 accumulate in TMEM); 50 launches rotating over 4 weight copies"})
            if mh:
                L_.rwkv_b200_matrix_free(ctypes.c_void_p(mh))
        except Exception as ex:                          # the roofline leg never fails the bench line
            roof["error"] = str(ex)
        out = {"metric": f"prefill tokens/s ({shape} Q5_1, T={T})", "value": d.world * T / (ms_max * 1e-3), "unit": "tokens/s", "n_gpus": d.world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
               "dtype": "f16 operands (weights dequantised in-kernel), f32 accumulate (tcgen05 / TMEM); f32 recurrence", "data": "synthetic (random-init weights in ggmf format, uniform random token ids)",
               "config": {"workload": f"{shape} Q5_1 rwkv_eval_sequence T={T} (BASELINE.json configs[2])", "algorithmic_flops_per_step": flops,
                          "l2": "weights 1.26 GB per pass > L2"},
               "clocks": clocks, "e2e": {"value": d.world * T / e2e, "unit": "tokens/s", "h2d_bytes_per_step": T * 4, "d2h_bytes_per_step": (model.state_len + model.n_vocab) * 4,
                                         "api": "rwkv_eval_sequence(ctx, tokens, T, NULL, state_out, logits_out) with host buffers"},
               "gpu_launches": (12 * L + 3) * args.steps if sh.arch == 6 else None,      # v6 sequence pass: 12 launches per layer + embed, ln_out, head (profiles/ launch list)
               "roofline": roof}
        if not args.no_cpu:
            try:
                Lr = reference_lib()
                Lr.lib.rwkv_set_print_errors(None, False)
                n_thr = min(host_cores(), 32)
                ctx = Lr.load(path, n_thr, 0)
                Ts = 64                                   # bounded sample of the same workload: a 64-token sequence
                ctx.eval_sequence([int(t) for t in toks[:8]], None)
                t0 = time.perf_counter()
                ctx.eval_sequence([int(t) for t in toks[:Ts]], None)
                dt = time.perf_counter() - t0
                ctx.free()
                out["cpu_baseline"] = {"value": Ts / dt, "unit": "tokens/s", "cores": n_thr, "kind": "reference",
                                       "sample": f"one rwkv_eval_sequence of {Ts} tokens of the unmodified reference CPU build on the same model file; " + Lr.system_info()}
            except Exception as ex:
                out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": 0, "kind": "reference", "sample": "not run: " + str(ex)}
    model.free(); d.close()
    return out


def run_streams(args):
    """BASELINE configs[3]: v7 2.9B-shape Q8_0, 256 independent streams sharded over the GPUs."""
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    d = Dist()
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    lib.library.rwkv_b200_set_device(d.local)
    shape = args.shape or "v7-2b9"
    if d.rank == 0:
        synth.ensure_model(lib, shape, "Q8_0")
    d.barrier()
    path = synth.ensure_model(lib, shape, "Q8_0")
    model = rwkv_cppy_b200.RWKVModel(lib, path)
    B = args.streams // d.world
    batch = model.create_batch(B)
    K, W = args.steps, max(args.warmup, 3)
    toks = np.random.default_rng(1 + d.rank).integers(0, model.n_vocab, size=(W + K, B)).astype(np.uint32)
    batch.run_timed(toks[:W], True)
    d.barrier()
    sampler = ClockSampler(d.local); sampler.start()
    ms = batch.run_timed(toks[W:], True)
    d.barrier()
    clocks = sampler.stop()
    ms_max = d.max(ms)
    t0 = time.perf_counter()
    for s in range(min(K, 8)):
        batch.eval(toks[W + s])
    e2e = (time.perf_counter() - t0) / min(K, 8)
    launches = batch.launches_per_step(True)
    out = None
    if d.rank == 0:
        # algorithmic bytes per step per GPU: every weight byte the step multiplies once (file - embedding + B embedding rows)
        # + each stream's state read and written
        sh = synth.SHAPES[shape]
        wbytes = os.path.getsize(path) - sh.n_vocab * sh.n_embed * 2 + B * sh.n_embed * 2
        step_bytes = wbytes + 2 * B * model.state_len * 4
        hbm, _, which = peaks()
        out = {"metric": f"decode tokens/s ({shape} Q8_0, {args.streams} independent streams)", "value": d.world * B * K / (ms_max * 1e-3), "unit": "tokens/s",
               "n_gpus": d.world, "steps": K, "warmup": W, "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "int8 dp4a blocks + f32 accumulate", "data": "synthetic", "config": {"workload": f"{shape} Q8_0, {B} streams per GPU, state resident (BASELINE.json configs[3])"},
               "clocks": clocks, "gpu_launches": launches * K,
               "e2e": {"value": d.world * B / e2e, "unit": "tokens/s", "h2d_bytes_per_step": B * 4, "d2h_bytes_per_step": B * model.n_vocab * 4,
                       "api": "rwkv_b200_batch_eval (host tokens in, host logits out)"},
               "roofline": {"bound": "hbm", "achieved": step_bytes / (ms_max / K * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                            "frac": step_bytes / (ms_max / K * 1e-3) / 1e9 / hbm, "traffic": None,
                            "what": "whole batch step per GPU: weights once + every stream's state read and written", "algorithmic_bytes_per_gpu_per_step": step_bytes,
                            "peak_source": which}}
        if not args.no_cpu:
            try:                                           # one stream of the same workload on the host cores (the reference has no batch API: streams are a loop over rwkv_eval)
                tok_s, thr, info, _, runs = time_reference_decode(path, toks[:, 0], 2, 6, repeats=3)
                out["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference",
                                       "sample": "best of 3 runs of 6 rwkv_eval calls of ONE stream (the reference evaluates streams one after the other) on the same model file; " + info}
            except Exception as ex:
                out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": 0, "kind": "reference", "sample": "not run: " + str(ex)}
    batch.free(); model.free(); d.close()
    return out


def run_tp(args):
    """BASELINE configs[4]: v4 14B-shape, single-stream decode tensor-parallel over the ranks (NCCL)."""
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    d = Dist()
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    lib.library.rwkv_b200_set_device(d.local)
    shape = args.shape or "v4-14b"
    fmt = args.format or "FP16-to-Q4_1"
    if d.rank == 0:
        synth.ensure_model(lib, shape, fmt)
    d.barrier()
    path = synth.ensure_model(lib, shape, fmt)
    if d.world > 1:
        ids = [lib.rwkv_b200_tp_unique_id() if d.rank == 0 else None]
        d.dist.broadcast_object_list(ids, src=0)
        model = rwkv_cppy_b200.RWKVModel(lib, path, tensor_parallel=(d.rank, d.world, ids[0]))
        if os.environ.get("RWKV_B200_TP_NCCL_ONLY") != "1":      # one-shot all-reduce over NVLink peer memory instead of NCCL
            handles = [None] * d.world
            d.dist.all_gather_object(handles, lib.rwkv_b200_tp_ipc_handle(model._ctx))
            lib.rwkv_b200_tp_connect(model._ctx, handles)
    else:
        model = rwkv_cppy_b200.RWKVModel(lib, path)
    batch = model.create_batch(1)
    K, W = args.steps, max(args.warmup, 3)
    toks = np.random.default_rng(1).integers(0, model.n_vocab, size=(W + K, 1)).astype(np.uint32)     # same stream on every rank
    batch.run_timed(toks[:W], True)
    d.barrier()
    sampler = ClockSampler(d.local); sampler.start()
    ms = batch.run_timed(toks[W:], True)
    d.barrier()
    clocks = sampler.stop()
    ms_max = d.max(ms)
    # end to end through the C ABI: host state in / out and logits out on every rank
    state = np.zeros(model.state_len, np.float32); logits = np.zeros(model.n_vocab, np.float32)
    n_e2e = min(K, 16)
    d.barrier()
    t0 = time.perf_counter()
    for i in range(n_e2e):
        assert lib.library.rwkv_eval(model._ctx.ptr, int(toks[W + i, 0]), state.ctypes.data_as(PF), state.ctypes.data_as(PF), logits.ctypes.data_as(PF))
    e2e = d.max((time.perf_counter() - t0) / n_e2e)
    launches = batch.launches_per_step(True)
    shp = synth.SHAPES[shape]
    wbytes = os.path.getsize(path)
    out = None
    if d.rank == 0:
        step_ms = ms_max / K
        out = {"metric": f"decode tokens/s ({shape} {fmt}, 1 stream, tensor parallel over {d.world} GPU)", "value": K / (ms_max * 1e-3), "unit": "tokens/s",
               "n_gpus": d.world, "steps": K, "warmup": W, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "int8 dp4a blocks + f32 accumulate" if "Q" in fmt else "f16 weights, f32 accumulate", "data": "synthetic",
               "config": {"workload": f"{shape} {fmt} single-stream decode, TP={d.world} (column/row-split weights, all-reduce x2 per layer + logits all-gather)",
                          "collective": "NCCL" if os.environ.get("RWKV_B200_TP_NCCL_ONLY") == "1" or d.world == 1 else "one-shot P2P kernel over NVLink peer memory (csrc/tp_p2p.cu)",
                          "file_bytes": wbytes, "bytes_per_gpu_per_step": wbytes / d.world,
                          "hbm_gbs_per_gpu": wbytes / d.world / (step_ms * 1e-3) / 1e9, "layers": shp.n_layer},
               "clocks": clocks, "gpu_launches": launches * K,
               "e2e": {"value": 1.0 / e2e, "unit": "tokens/s", "h2d_bytes_per_step": model.state_len * 4 + 4, "d2h_bytes_per_step": (model.state_len + model.n_vocab) * 4,
                       "api": "rwkv_eval on every rank (host state in/out, full logits out)"}}
    batch.free(); model.free(); d.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="main", choices=["main", "decode", "matvec", "gemv", "prefill", "streams", "tp"])
    ap.add_argument("--no-extras", action="store_true", help="main workload: skip the records of the other BASELINE configs")
    ap.add_argument("--shape", default=None)
    ap.add_argument("--format", default=None, help="tp workload: model format (FP16 or FP16-to-Q4_1)")
    ap.add_argument("--tokens", type=int, default=1024)
    ap.add_argument("--streams", type=int, default=256)
    ap.add_argument("--only", type=int, default=None, help="gemv workload: index of the single shape to run")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        out = run_reference(args)
    else:
        out = {"main": run_main, "decode": run_decode, "matvec": run_gemv, "gemv": run_gemv, "prefill": run_prefill, "streams": run_streams, "tp": run_tp}[args.workload](args)
    Dist.shutdown()
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""bench.py -- throughput of the hot path on B200 (contract: see the task statement).

Default workload = BASELINE.json configs[1]: RWKV v4 Pile-169M shape, Q4_0 body + FP16
head, random-init weights written in the ggmf format, single-token decode of one stream per
GPU with logits.  One "step" = one token through every layer + head.

  value  tokens/s with state + token ids resident in HBM (CUDA-graph replay, device-timed)
  e2e    tokens/s through the rwkv.h C ABI (rwkv_eval with HOST state/logits buffers,
         host<->device copies inside the timed region)
  roofline       the dominant kernel of the step (the FP16 head mat-vec, 77 MB) timed with
                 CUDA events on the engine's stream over rotating copies (> L2)
  cpu_baseline   the unmodified reference CPU library (oracle/_ref) on the same model file
  --impl reference  times that reference alone (all host threads it can use)

Other workloads (`--workload`) report the remaining BASELINE configs the same way:
  gemv      the Q5_1 mat-vec microbenchmark (HBM GB/s on 14B / 1.6B shapes)
  prefill   v6 1.6B-shape Q5_1, one 1024-token rwkv_eval_sequence
  streams   v7 2.9B-shape Q8_0, 256 independent streams sharded over the GPUs
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


def time_reference_decode(model_file: str, tokens, warmup: int, steps: int, threads=None):
    """tokens/s of the reference's rwkv_eval loop on host cores; picks the best thread count
    from a short sweep when `threads` is None.  Returns (tok_s, threads, sysinfo)."""
    L = reference_lib()
    L.lib.rwkv_set_print_errors(None, False)
    cores = host_cores()
    cands = [threads] if threads else sorted({c for c in (4, 8, 16, 32, cores // 2, cores) if 1 <= c <= cores})
    best = (0.0, cands[0])
    for n in cands:
        ctx = L.load(model_file, n, 0)
        st = None
        lg = np.empty(ctx.logits_len, np.float32)
        for t in tokens[:max(2, warmup // 2)]:
            st, _ = ctx.eval(int(t), st, state_out=st, logits_out=lg)
        t0 = time.perf_counter()
        k = max(4, min(steps, 16)) if len(cands) > 1 else steps
        for t in tokens[:k]:
            st, _ = ctx.eval(int(t), st, state_out=st, logits_out=lg)
        dt = time.perf_counter() - t0
        ctx.free()
        if k / dt > best[0]:
            best = (k / dt, n)
    if len(cands) > 1:                       # full-length run at the best thread count
        ctx = L.load(model_file, best[1], 0)
        st = None
        lg = np.empty(ctx.logits_len, np.float32)
        for t in tokens[:warmup]:
            st, _ = ctx.eval(int(t), st, state_out=st, logits_out=lg)
        t0 = time.perf_counter()
        for t in tokens[:steps]:
            st, _ = ctx.eval(int(t), st, state_out=st, logits_out=lg)
        dt = time.perf_counter() - t0
        ctx.free()
        best = (steps / dt, best[1])
    return best[0], best[1], L.system_info()


class Dist:
    """torch.distributed only as plumbing: barrier + max over ranks."""

    def __init__(self, use_cuda: bool = True):
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
        if self.world > 1:
            self.dist.destroy_process_group()


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
            "config": dict(DECODE_CFG, l2="256 MiB scratch written before every timed step (cold L2), one CUDA event pair per step",
                           value_l2_warm=d.world * K / (ms_warm_max * 1e-3), algorithmic_bytes_per_step=129.8e6,
                           step_hbm_gbs=129.8e6 / (ms_max / K * 1e-3) / 1e9),
            "clocks": clocks,
            "e2e": {"value": d.world * K / e2e_s, "unit": "tokens/s", "h2d_bytes_per_step": S * 4 + 4, "d2h_bytes_per_step": S * 4 + V * 4,
                    "api": "rwkv_eval(ctx, token, state_in, state_out, logits_out) with pinned host buffers"},
            "gpu_launches": launches * K,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": ncu_traffic("head_gemv_f16_50277x768"),
                         "kernel": "gemv_kernel<F16,1> head 50277x768 (77.2 MB), dominant launch of the step",
                         "us_per_launch": kms * 1e3, "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs)" if which == "measured" else which},
        }
        if not args.no_cpu:
            try:
                tok_s, thr, info = time_reference_decode(path, toks[:, 0], 4, 48)
                out["cpu_baseline"] = {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference",
                                       "sample": "48 rwkv_eval calls (after 4 warm-up) of the unmodified reference CPU build on the same model file; " + info}
            except Exception as e:  # the baseline is a report, not a dependency
                out["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": 0, "kind": "reference", "sample": f"not run: {e}"}
    model.free()
    d.close()
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the same workload."""
    rank, world, _ = dist_env()
    if rank != 0:
        return None
    import rwkv_cppy_b200
    from rwkv_cppy_b200 import synth
    lib = rwkv_cppy_b200.load_rwkv_shared_library()            # only its host-side quantiser is used here
    path = synth.ensure_model(lib, DECODE_CFG["shape"], DECODE_CFG["format"])
    K, W = args.steps, max(args.warmup, 3)
    K = min(K, 96)                                             # bounded sample: ~10 ms/token on host cores
    rng = np.random.default_rng(1)
    toks = rng.integers(0, 50277, size=W + K).astype(np.uint32)
    tok_s, thr, info = time_reference_decode(path, toks, W, K)
    return {"impl": "reference", "metric": "decode tokens/s (RWKV v4 169M-shape Q4_0, single stream per GPU)", "value": tok_s, "unit": "tokens/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": 1e3 / tok_s, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int8 blocks + f32 accumulate (ggml CPU)", "data": "synthetic (same model file and token stream)",
            "config": DECODE_CFG,
            "cpu_baseline": {"value": tok_s, "unit": "tokens/s", "cores": thr, "kind": "reference",
                             "sample": f"{K} rwkv_eval calls of the unmodified reference (oracle/_ref) on rank 0; best of a thread sweep; " + info},
            "e2e": {"value": tok_s, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


GEMV_SHAPES = [("Q5_1", 8, 5120, 5120), ("Q5_1", 8, 20480, 5120), ("Q5_1", 8, 5120, 20480), ("Q5_1", 8, 7168, 2048),
               ("Q4_0", 2, 20480, 5120), ("Q8_0", 9, 20480, 5120), ("FP16", 1, 65536, 2048), ("FP16", 1, 50277, 768)]


def run_gemv(args):
    """BASELINE headline kernel metric: achieved HBM GB/s of the dequant-fused mat-vec."""
    import rwkv_cppy_b200
    from oracle import rwkv_oracle as O          # only to produce valid ggml blocks for the synthetic matrices
    lib = rwkv_cppy_b200.load_rwkv_shared_library()
    hbm, _, which = peaks()
    rng = np.random.default_rng(0)
    rows = []
    shapes = GEMV_SHAPES if args.only is None else [GEMV_SHAPES[args.only]]
    for name, tid, M, K in shapes:
        w = (rng.random((min(M, 4096), K), dtype=np.float32) * 2 - 1) / np.float32(np.sqrt(K))
        raw1 = w.astype(np.float16).view(np.uint8).reshape(-1) if tid == 1 else O.quantize_ref(tid, w.reshape(-1))
        raw = np.tile(raw1, (M + w.shape[0] - 1) // w.shape[0])[: (raw1.size // w.shape[0]) * M]
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
        synth.ensure_model(lib, shape, "FP16-to-Q5_1")
    d.barrier()
    path = synth.ensure_model(lib, shape, "FP16-to-Q5_1")
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
            from oracle import rwkv_oracle as O          # only to produce valid ggml blocks for a synthetic matrix
            rngw = np.random.default_rng(0)
            w = (rngw.random((256, D), dtype=np.float32) * 2 - 1) / np.float32(np.sqrt(D))
            raw1 = O.quantize_ref(8, w.reshape(-1))
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
               "config": {"workload": f"{shape} FP16-to-Q5_1 rwkv_eval_sequence T={T} (BASELINE.json configs[2])", "algorithmic_flops_per_step": flops,
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
        synth.ensure_model(lib, shape, "FP16-to-Q8_0")
    d.barrier()
    path = synth.ensure_model(lib, shape, "FP16-to-Q8_0")
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
        out = {"metric": f"decode tokens/s ({shape} Q8_0, {args.streams} independent streams)", "value": d.world * B * K / (ms_max * 1e-3), "unit": "tokens/s",
               "n_gpus": d.world, "steps": K, "warmup": W, "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "int8 dp4a blocks + f32 accumulate", "data": "synthetic", "config": {"workload": f"{shape} FP16-to-Q8_0, {B} streams per GPU, state resident"},
               "clocks": clocks, "gpu_launches": launches * K,
               "e2e": {"value": d.world * B / e2e, "unit": "tokens/s", "h2d_bytes_per_step": B * 4, "d2h_bytes_per_step": B * model.n_vocab * 4,
                       "api": "rwkv_b200_batch_eval (host tokens in, host logits out)"}}
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
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="decode", choices=["decode", "gemv", "prefill", "streams", "tp"])
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
        out = {"decode": run_decode, "gemv": run_gemv, "prefill": run_prefill, "streams": run_streams, "tp": run_tp}[args.workload](args)
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

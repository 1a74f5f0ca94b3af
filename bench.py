#!/usr/bin/env python
"""bench.py — the reference's headline metric on its headline config (BASELINE.json):
Mistral-7B Q4_K_M, batch-1 greedy decode, 512-token context (384 prompt + 128 generated), tok/s.

One "step" = one 128-token greedy decode from position 384 to 512 on a seeded random-init GGUF of
the Mistral-7B architecture (synthetic data; there is no network for real checkpoints).  The
4.29 GB of weights streamed per token exceed the 126 MB L2, so no L2 flush is needed between steps.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA engine
  python bench.py --impl reference [...]                         # the CPU path, timed on host cores

Multi-GPU (torchrun, one rank per GPU): the path shards by independent requests, so every rank
runs its own full replica (SURVEY.md §8e "replicas only"), no collective on the data path;
value = tokens of all ranks / max-over-ranks time ("weak" scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")  # the oracle's OpenMP team must not spin against other busy threads

import numpy as np  # noqa: E402

from aios_b200 import gguf_synth as gs  # noqa: E402

SEED = 0xA105
METRIC = "Mistral-7B Q4_K_M decode tok/s (bs1, ctx512)"


def model_path(arch: str, ftype: str) -> str:
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
    return os.path.join(base, f"aios_bench_{arch}_{ftype}_{SEED:x}.gguf")


def ensure_model(arch: str, ftype: str, writer: bool) -> str:
    path = model_path(arch, ftype)
    want = None
    if writer and not os.path.exists(path):
        tmp = path + f".tmp{os.getpid()}"
        info = gs.write_synthetic_gguf(tmp, arch, ftype, SEED)
        os.replace(tmp, path)
        want = info.file_bytes
    return path


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines: list[str] = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def recorded_traffic() -> float | None:
    """dram bytes per token of the GEMV launches from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return float(json.load(open(p))["gemv_dram_bytes_per_token"])
    except Exception:
        return None


def host_threads() -> int:
    """Threads the CPU arm may use: the cores this process is allowed on.  (torchrun injects OMP_NUM_THREADS=1 into its
    ranks, which would silently time the CPU path on one core; the oracle's pool is sized explicitly instead.)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def oracle_window(path: str, arch: str, n_prompt: int, n_gen: int):
    """CPU restatement of the llama.cpp CPU path (oracle, AVX2 integer dots, all host threads) on the bench's own
    window: feeds the n_prompt-token prompt position by position, then decodes n_gen greedy tokens.  Returns
    (tokens, forward passes per second, threads, seconds, forward passes)."""
    from oracle import oracle as orc

    orc.build()
    threads = host_threads()
    m = orc.OracleModel(path, n_ctx=n_prompt + n_gen + 8, simd=True, threads=threads)
    prompt = gs.synthetic_prompt(arch, n_prompt, SEED)
    t0 = time.perf_counter()
    toks, _ = m.greedy(prompt, n_gen)
    dt = time.perf_counter() - t0
    m.close()
    passes = n_prompt + n_gen - 1  # greedy() does not feed the last token it returns
    return toks, passes / dt, orc.lib().oracle_num_threads(), dt, passes


def run_reference(args, rank: int, world: int):
    """The reference arm: the reference's CPU path for this workload.  llama.cpp itself is an un-vendored, un-pinned
    dependency of the reference and is not on this box, so this times the oracle (CPU restatement, AVX2 dots) with all
    the host threads it can use.  Same GGUF, same 384-token prompt (fed untimed, like the GPU arm's prefill); a step
    is a bounded sample of the GPU arm's step: the first `--ref-tokens-per-step` greedy tokens of the 128-token
    window, restarted at position 384 every step."""
    if rank != 0:
        return
    path = ensure_model(args.arch, args.ftype, True)
    from oracle import oracle as orc

    orc.build()
    threads = host_threads()
    n_tok = min(args.ref_tokens_per_step, args.gen)
    m = orc.OracleModel(path, n_ctx=args.prompt + args.gen + 8, simd=True, threads=threads)
    cores = orc.lib().oracle_num_threads()
    prompt = gs.synthetic_prompt(args.arch, args.prompt, SEED)
    logits0 = m.eval(prompt)  # untimed: positions 0..prompt-1 (the GPU arm's prefill is untimed too)
    first = int(np.argmax(logits0))

    def step():
        tok = np.array([first], np.int32)
        pos = args.prompt
        for _ in range(n_tok):  # decode step: feed the token at `pos`, arg-max the next one
            lg = m.eval(tok, n_past=pos)  # rows >= prompt of the KV cache are simply overwritten each step
            tok = np.array([int(np.argmax(lg))], np.int32)
            pos += 1

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = args.steps * n_tok / dt
    sample = (f"{args.steps} steps x the first {n_tok} of the {args.gen} greedy decode tokens at context {args.prompt}.."
              f"{args.prompt + n_tok} (same GGUF, same {args.prompt}-token prompt fed untimed); CPU restatement with AVX2 "
              f"integer dots on {cores} threads, not llama.cpp (un-vendored by the reference, unavailable on this box)")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int8 x q4/q6 integer dots, f32/f64 accumulate", "data": "synthetic",
        "config": {"workload": f"{args.arch} {args.ftype} batch-1 greedy decode, {args.prompt}-token prompt + "
                               f"{args.gen} generated = {args.prompt + args.gen}-token context (BASELINE.json configs[1])",
                   "tokens_per_step": n_tok, "window": f"positions {args.prompt}..{args.prompt + n_tok}", "seed": SEED},
        "cpu_baseline": {"value": val, "unit": "tok/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def bench_config4(eng, model, arch: str, ftype: str, peak: float, n_lanes: int = 64, n_prompt: int = 128, n_gen: int = 128):
    """BASELINE.json configs[3]: batched decode, 64 concurrent requests on one GPU (one token per request per step, all
    through one weight stream).  Device-timed steps (CUDA events, every step's tokens copied to the host like in the
    scheduler) plus the same load end to end through aios_request_submit / poll (prefill included)."""
    info = model.info
    rng_prompts = np.stack([gs.synthetic_prompt(arch, n_prompt, SEED + 100 + i) for i in range(n_lanes)])
    model.bench_batch_decode(rng_prompts[:, :8], 4, want_tokens=False)  # warm-up: buffers, graphs of this lane count
    r, toks = model.bench_batch_decode(rng_prompts, n_gen)
    # parity spot check: two lanes against the single-sequence path
    for lane in (0, n_lanes - 1):
        alone = model.generate(rng_prompts[lane], n_gen)
        if not np.array_equal(alone, toks[lane]):
            raise SystemExit(f"bench.py: PARITY FAILURE — batched lane {lane} differs from the single-sequence path")
    steps_s = r.n_steps_timed / (r.ms_decode * 1e-3)
    tok_s = steps_s * n_lanes
    n_past_mean = (r.n_past_begin + r.n_past_end) / 2.0
    bytes_step = info.weight_bytes_per_token + n_lanes * info.kv_bytes_per_position * n_past_mean
    # end to end: submit, wait, read every token back through the request API (host buffers in, host buffers out)
    t0 = time.perf_counter()
    reqs = [model.submit(rng_prompts[i], n_gen) for i in range(n_lanes)]
    got = [q.result(600000) for q in reqs]
    dt = time.perf_counter() - t0
    for q in reqs:
        q.free()
    if not all(np.array_equal(g, t) for g, t in zip(got, toks)):
        raise SystemExit("bench.py: PARITY FAILURE — queued requests differ from the batched decode hook")
    return {"workload": f"{arch} {ftype} batched greedy decode, {n_lanes} concurrent requests x ({n_prompt}-token prompt + {n_gen} "
                        "generated) on one GPU (BASELINE.json configs[3])",
            "value": tok_s, "unit": "tok/s", "ms_per_step": r.ms_decode / r.n_steps_timed, "steps_timed": r.n_steps_timed,
            "lanes": n_lanes, "gpu_launches": r.kernel_launches, "bytes_per_step": bytes_step,
            "roofline_tok_s": n_lanes * peak * 1e9 / bytes_step, "frac_of_hbm_roofline": tok_s / (n_lanes * peak * 1e9 / bytes_step),
            "prefill_ms": r.ms_prefill,
            "e2e": {"value": n_lanes * n_gen / dt, "unit": "tok/s generated, prompt prefill included in the time",
                    "seconds": dt, "api": "aios_request_submit x64 + aios_request_wait / poll",
                    "h2d_bytes_per_step": int(rng_prompts.nbytes), "d2h_bytes_per_step": n_lanes * n_gen * 4},
            "parity_checked": True}


def measured_tensor_peak() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    except Exception:
        return 1400.0, "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)"


def bench_config3(device: int, n_prompt: int = 4096, iters: int = 3):
    """BASELINE.json configs[2]: Mistral-7B Q8_0, prefill of a 4096-token prompt through the tensor-core path (tcgen05 /
    TMEM dequant GEMM + causal flash attention).  tok/s and achieved / measured-sustained bf16 TFLOP/s (SURVEY.md §8d:
    2 * 6.979e9 * T mat-mul + 4 * n_head * hd * T^2 / 2 causal attention FLOPs)."""
    from aios_b200 import capi
    arch, ftype = "mistral-7b", "Q8_0"
    path = ensure_model(arch, ftype, True)
    eng = capi.Engine(device=device, max_batch=1)  # default flags: prompts of 256+ tokens take the tensor-core path
    try:
        m = eng.load_model("bench-q8", path, n_ctx=n_prompt + 8)
        try:
            info = m.info
            a = gs.ARCHS[arch]
            prompt = gs.synthetic_prompt(arch, n_prompt, SEED + 3)
            s = m.new_seq()
            try:
                ms = s.bench_prefill(prompt, iters)
                # end to end through the C ABI: token ids from host memory in, the last token's logits back to the host
                s.reset()
                t0 = time.perf_counter()
                s.prefill(prompt)
                lg = s.logits()
                e2e_s = time.perf_counter() - t0
                finite = bool(np.isfinite(lg).all())
            finally:
                s.free()
            per_tok_mm = 2.0 * a.n_layer * (a.n_embd * (a.n_head * a.head_dim + 2 * a.n_head_kv * a.head_dim) + a.n_head * a.head_dim * a.n_embd + 3 * a.n_embd * a.n_ff)
            flop = per_tok_mm * n_prompt + 4.0 * a.n_layer * a.n_head * a.head_dim * n_prompt * n_prompt / 2.0
            peak, src = measured_tensor_peak()
            tflops = flop / (ms * 1e-3) / 1e12
            return {"workload": f"{arch} {ftype} prefill of a {n_prompt}-token prompt (BASELINE.json configs[2]), tensor-core path: "
                                "tcgen05 / TMEM dequant GEMM, TMA-staged f16 activations, causal flash attention",
                    "value": n_prompt / (ms * 1e-3), "unit": "tok/s", "ms_per_prefill": ms, "iters": iters,
                    "flop_per_prefill": flop, "roofline": {"bound": "tensor", "achieved": tflops, "peak": peak, "unit": "TFLOP/s",
                                                           "frac": tflops / peak, "peak_source": src},
                    "roofline_tok_s": n_prompt / (flop / (peak * 1e12)),
                    "e2e": {"value": n_prompt / e2e_s, "unit": "tok/s", "h2d_bytes_per_step": 4 * n_prompt,
                            "d2h_bytes_per_step": 4 * info.n_vocab, "api": "aios_seq_prefill (host token ids) + aios_seq_logits"},
                    "numerics": "f16 operands, f32 accumulation; stated tolerance in tests/test_gpu_prefill_tc.py", "logits_finite": finite}
        finally:
            m.unload()
    finally:
        eng.close()


def bench_served(model, arch: str, n_prompt: int = 256, n_gen: int = 128, reps: int = 3):
    """What one agent call costs: unary Infer semantics through the C ABI's request queue — JSON mode (the reference
    always asks for response_format json_object, runtime/src/inference.rs:119-121), temperature 0.3, a 256-token prompt,
    128 generated tokens.  The grammar runs in C++ on the scheduler thread, its mask goes to the GPU every step
    (n_vocab / 8 bytes H2D), the sampled token comes back (4 B D2H).  Decode rate = tokens after the first / time after
    the first token (the prompt's prefill is reported separately)."""
    from aios_b200 import capi
    vocab = model.vocab()
    prompt = gs.synthetic_prompt(arch, n_prompt, SEED + 9)
    rates, ttft, n_out = [], [], 0
    for rep in range(reps + 1):
        g = capi.JsonGrammar(vocab)
        t0 = time.perf_counter()
        r = model.submit(prompt, n_gen, capi.make_sampling(0.3, 40, 0.95, 0.05, seed=SEED + rep), json_grammar=g)
        t_first, n = None, 0
        while True:
            r.wait(60000)
            toks, done = r.poll(4096)
            if toks.size and t_first is None:
                t_first, n_first = time.perf_counter(), int(toks.size)
            n += int(toks.size)
            if done:
                break
        t1 = time.perf_counter()
        r.free()
        g.free()
        if rep and t_first is not None and n > n_first:
            rates.append((n - n_first) / (t1 - t_first))
            ttft.append(t_first - t0)
            n_out = n
    return {"workload": f"unary Infer as served: JSON mode (device logit mask), temperature 0.3, {n_prompt}-token prompt, up to {n_gen} "
                        "tokens, one request at a time through aios_request_submit_masked",
            "value": float(np.median(rates)) if rates else None, "unit": "tok/s", "tokens_generated": n_out,
            "seconds_to_first_token": float(np.median(ttft)) if ttft else None,
            "h2d_bytes_per_token": (model.info.n_vocab + 7) // 8, "d2h_bytes_per_token": 4}


def in_graph_breakdown(seq, n_runs: int = 4):
    """Per-token time of each launch kind INSIDE the token pass, from the %globaltimer marks the kernels themselves take
    (aios_seq_trace_decode: same launch order and programmatic-dependent-launch attributes as the captured graph).
    A launch's share is completion-to-completion: from the moment the previous launch's last CTA finished to the moment
    its own last CTA finished — hand-over, dependency wait, activation prologue and weight stream included — so the
    shares of all launches add up to the token's span and never exceed it.
    Returns dict(us_token, us_gemv, us_attn, us_other, n_gemv, n_attn)."""
    acc = {"us_token": 0.0, "us_gemv": 0.0, "us_attn": 0.0, "us_other": 0.0}
    n_gemv = n_attn = 0
    for _ in range(n_runs):
        seq.restore()
        tr = seq.trace_decode().astype(np.int64)
        done = np.maximum(tr[:, 7], tr[:, 4])
        kind = tr[:, 3]
        traced = tr[:, 0] > 0
        idx = np.nonzero(traced)[0]
        if idx.size < 2:
            continue
        prev = tr[idx[0], 0]  # the first traced launch's entry opens the token
        n_gemv = int((kind[idx] == 1).sum())
        n_attn = int((kind[idx] == 2).sum())
        for i in idx:
            d = max(0, int(done[i] - prev)) / 1e3
            prev = max(prev, int(done[i]))
            acc["us_gemv" if kind[i] == 1 else "us_attn" if kind[i] == 2 else "us_other"] += d
        acc["us_token"] += (prev - tr[idx[0], 0]) / 1e3
    out = {k: v / n_runs for k, v in acc.items()}
    out.update(n_gemv=n_gemv, n_attn=n_attn)
    return out


def run_tp(args):
    """BASELINE.json configs[4]: one model split over --tp GPUs of this node (one process drives every shard).

    value = device-timed greedy decode tok/s from position prompt to prompt + gen; e2e = the same tokens through
    aios_tp_prefill (last prompt token from host memory) + aios_tp_decode (tokens back to a host buffer).  The first
    greedy tokens are printed (`first_tokens`) so that runs at different TP degrees — and the single-GPU engine, where
    the model fits — can be compared: the split does not change the tokens (f64 partial sums, rounded once)."""
    import hashlib
    import shutil

    os.environ["NCCL_DEBUG"] = "WARN"  # NCCL's version banner goes to stdout, where the JSON line belongs

    from aios_b200 import capi

    n = args.tp
    arch, ftype = args.tp_arch, args.ftype
    a = gs.ARCHS[arch]
    path = model_path(arch, ftype)
    if not os.path.exists(path):
        approx = (a.n_layer * (2 * a.n_embd * a.n_head * a.head_dim + 2 * a.n_embd * a.n_head_kv * a.head_dim + 3 * a.n_embd * a.n_ff)
                  + 2 * a.n_vocab * a.n_embd) * 0.6
        free = shutil.disk_usage(os.path.dirname(path)).free
        if free < approx * 1.15:
            print(json.dumps({"metric": f"{arch} {ftype} decode tok/s (bs1, TP-{n})", "unavailable":
                              f"the synthetic GGUF needs ~{approx / 1e9:.0f} GB, {free / 1e9:.0f} GB free"}), flush=True)
            return
    path = ensure_model(arch, ftype, True)
    n_ctx = args.prompt + args.gen + 16  # bench_decode: 2 untimed tokens, the timed window, 8 tokens with per-all-reduce events
    t0 = time.perf_counter()
    tp = capi.TpModel(list(range(n)), "bench-tp", path, n_ctx=n_ctx)
    load_s = time.perf_counter() - t0
    info = tp.info
    prompt = gs.synthetic_prompt(arch, args.prompt, SEED)
    first = tp.generate(prompt[:32], 16)
    again = tp.generate(prompt[:32], 16)
    for _ in range(max(args.warmup, 3)):
        tp.reset()
        tp.prefill(prompt[:8])
        tp.bench_decode(8)
    peak, peak_src = measured_peaks()
    sampler = ClockSampler(0)
    sampler.start()
    dev_ms, launches, tokens, ar_us = 0.0, 0, 0, []
    kind = graphs = None
    for _ in range(args.steps):
        tp.reset()
        tp.prefill(prompt)  # untimed; token by token through the decode pass (no tensor-parallel prefill GEMM yet)
        r = tp.bench_decode(args.gen)
        dev_ms += r.ms_total
        launches += r.kernel_launches
        tokens += r.n_tokens
        ar_us.append(r.us_per_allreduce)
        kind, graphs = r.allreduce_kind, r.graphs
    clocks = sampler.stop()
    last_tok = np.ascontiguousarray(prompt[-1:], dtype=np.int32)
    e2e_tokens, e2e_ms = 0, 0.0
    for _ in range(args.steps):
        tp.reset()
        tp.prefill(prompt[:-1])
        t0 = time.perf_counter()
        tp.prefill(last_tok)
        e2e_tokens += tp.decode(args.gen).size
        e2e_ms += (time.perf_counter() - t0) * 1e3
    value = tokens / (dev_ms * 1e-3)
    ms_token = dev_ms / tokens
    rank_bytes = info.weight_bytes_per_token  # one shard's algorithmic weight bytes per token
    line = {
        "metric": f"{arch} {ftype} decode tok/s (bs1, TP-{n})", "value": value, "unit": "tok/s", "n_gpus": n, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "int8 x q4/q5/q6 integer dots, f32 terms summed in f64 across ranks", "data": "synthetic",
        "config": {"workload": f"{arch} {ftype} batch-1 greedy decode split {n} ways (BASELINE.json configs[4]), "
                               f"{args.gen} tokens per step from position {args.prompt + 2}", "parallelism": f"tp{n}, one process",
                   "l2": f"inputs larger than L2 ({rank_bytes / 1e9:.1f} GB of weights per rank per token)",
                   "allreduce": "partial sums pushed from the GEMV epilogue over NVLink peer mappings + arrival flags"
                                if kind else "ncclAllReduce (f64) per projection",
                   "cuda_graph_per_rank": bool(graphs), "allreduces_per_token": 2 * info.n_layer,
                   "us_per_allreduce": float(np.mean(ar_us)), "load_s": round(load_s, 1), "seed": SEED},
        "roofline": {"bound": "hbm", "achieved": rank_bytes / (ms_token * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": rank_bytes / (ms_token * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                     "method": "one rank's algorithmic weight bytes per token / device time per token (every rank streams "
                               "its shard concurrently; attention, all-reduce waits and the sampler are inside)",
                     "algorithmic_bytes_per_token_per_rank": rank_bytes},
        "e2e": {"value": e2e_tokens / (e2e_ms * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 4 * args.gen,
                "api": "aios_tp_prefill(last prompt token, host memory) + aios_tp_decode(gen tokens to a host buffer)"},
        "gpu_launches": launches, "clocks": clocks, "repeatable": bool(np.array_equal(first, again)),
        "first_tokens": [int(t) for t in first], "first_tokens_sha1": hashlib.sha1(first.tobytes()).hexdigest()[:12],
    }
    print(json.dumps(line), flush=True)
    tp.unload()
    if not args.keep_model and os.path.getsize(path) > 8e9:
        os.remove(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--arch", default="mistral-7b")
    ap.add_argument("--ftype", default="Q4_K_M")
    ap.add_argument("--prompt", type=int, default=384)
    ap.add_argument("--gen", type=int, default=128)
    ap.add_argument("--ref-tokens-per-step", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the oracle run (parity gate + cpu_baseline)")
    ap.add_argument("--parity-tokens", type=int, default=16)
    ap.add_argument("--no-subrecords", action="store_true", help="headline measurement only (no config3 / config4 sub-records)")
    ap.add_argument("--tp", type=int, default=0, help="tensor-parallel decode of --tp-arch over this many GPUs (one process)")
    ap.add_argument("--tp-arch", default="llama3-70b")
    ap.add_argument("--keep-model", action="store_true", help="leave the large synthetic GGUF in place for the next run")
    args = ap.parse_args()
    if args.tp:
        run_tp(args)
        return
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"  # NCCL's version banner goes to stdout, where the one JSON line belongs
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from aios_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this engine has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    path = ensure_model(args.arch, args.ftype, writer=(local_rank == 0))
    barrier()
    n_ctx = args.prompt + args.gen
    # seq, seq_e2e, and a temporary sequence for the parity spot checks; the 384-token prompt goes through the bit-exact
    # integer prefill so that the decode that follows can be checked against the oracle token by token
    eng = capi.Engine(device=local_rank, max_batch=3, flags=capi.ENGINE_EXACT_PREFILL)
    model = eng.load_model("bench", path, n_ctx=n_ctx)
    info = model.info
    seq = model.new_seq()
    prompt = gs.synthetic_prompt(args.arch, args.prompt, SEED)
    seq.prefill(prompt)
    seq.snapshot()
    # for the end-to-end leg: the same prompt minus its last token, which every step feeds from host memory
    seq_e2e = model.new_seq()
    seq_e2e.prefill(prompt[:-1])
    seq_e2e.snapshot()
    last_tok = np.ascontiguousarray(prompt[-1:], dtype=np.int32)

    # parity gate, before anything is timed: the first tokens of this very run against the oracle on the same GGUF and
    # prompt (rank 0; the same CPU run is the cpu_baseline sample).  A mismatch aborts the bench.
    parity = None
    cpu = None
    # (single-GPU runs only: under torchrun the other ranks would spin in the barrier on the cores the oracle needs;
    # every rank runs the same code on the same file, and the driver's N = 1 run gates it)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_chk = min(args.parity_tokens, args.gen)
        want, cpu_rate, cores, cpu_dt, cpu_passes = oracle_window(path, args.arch, args.prompt, n_chk)
        got = seq.decode(n_chk)
        seq.restore()
        if not np.array_equal(got, want):
            raise SystemExit(f"bench.py: PARITY FAILURE — first {n_chk} greedy tokens differ from the oracle: {got} vs {want}")
        parity = n_chk
        cpu = {"value": cpu_rate, "unit": "tok/s", "cores": cores, "kind": "port",
               "sample": f"{cpu_passes} single-token forward passes on the same GGUF: the {args.prompt} prompt positions and the "
                         f"first {n_chk - 1} generated tokens, contexts 1..{args.prompt + n_chk - 1} ({cpu_dt:.1f} s of CPU work); "
                         "CPU restatement with AVX2 integer dots, not llama.cpp (unavailable on this box)"}
    barrier()

    for _ in range(max(args.warmup, 3)):
        seq.restore()
        seq.bench_decode(args.gen)

    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t_wall0 = time.perf_counter()
    dev_ms, launches = 0.0, 0
    for _ in range(args.steps):
        seq.restore()
        r = seq.bench_decode(args.gen)
        dev_ms += r.ms_total
        launches += r.kernel_launches
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop()

    # end to end through the C ABI: every token comes back to a host buffer (4 B D2H) before the next is asked for
    # (the step's input — the token that starts the generation — is copied from host memory inside the timed region)
    for _ in range(2):
        seq_e2e.restore()
        seq_e2e.prefill(last_tok)
        seq_e2e.step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        seq_e2e.restore()
        seq_e2e.prefill(last_tok)
        for _ in range(args.gen):
            seq_e2e.step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3

    t = torch.tensor([dev_ms, e2e_ms, wall_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max, wall_ms_max = [float(v) for v in t.tolist()]

    # roofline legs.  (1) in-graph: the kernels' own %globaltimer marks inside the token pass; (2) isolated: a CUDA-event
    # pair around every launch of a few eagerly launched tokens (stand-alone rate, ~3 us of event overhead per launch).
    ig = in_graph_breakdown(seq)
    seq.restore()
    prof = seq.profile_decode(16)
    n_gemv = int(prof.reserved[2])
    emb_row = gs.tensor_nbytes((1, info.n_embd), gs.Q4_K if args.ftype == "Q4_K_M" else gs.Q8_0)
    gemv_bytes_per_token = info.weight_bytes_per_token - emb_row  # matrices + the norm vectors their prologues read
    peak, peak_src = measured_peaks()
    n_past_mean = args.prompt + (args.gen - 1) / 2.0
    bytes_per_token = info.weight_bytes_per_token + info.kv_bytes_per_position * n_past_mean
    total_tokens = world * args.steps * args.gen
    value = total_tokens / (dev_ms_max * 1e-3)
    e2e_value = total_tokens / (e2e_ms_max * 1e-3)
    us_token = dev_ms_max * 1e3 / (args.steps * args.gen)
    gemv_gbs = gemv_bytes_per_token / (ig["us_gemv"] * 1e-6) / 1e9 if ig["us_gemv"] > 0 else 0.0
    iso_gbs = gemv_bytes_per_token * prof.n_tokens / (prof.ms_gemv * 1e-3) / 1e9

    config4 = served = None
    if rank == 0 and world == 1 and not args.no_subrecords:
        served = bench_served(model, args.arch)
        served["fraction_of_greedy_value"] = served["value"] / value if served.get("value") else None
        config4 = {"mistral-7b": bench_config4(eng, model, args.arch, args.ftype, peak)}
        tl_path = ensure_model("tinyllama-1.1b", "Q4_K_M", True)
        tl = eng.load_model("bench-tl", tl_path, n_ctx=512)
        try:
            config4["tinyllama-1.1b"] = bench_config4(eng, tl, "tinyllama-1.1b", "Q4_K_M", peak)
        finally:
            tl.unload()

    config3 = None
    if rank == 0 and world == 1 and not args.no_subrecords:
        seq.free(); seq_e2e.free(); model.unload(); eng.close()
        seq = seq_e2e = model = eng = None
        config3 = bench_config3(local_rank)

    if rank == 0:
        # the dominant kernel's in-graph time must fit inside the token it is part of
        fits = ig["us_gemv"] <= us_token * 1.05 and ig["us_token"] <= us_token * 1.25
        traffic = recorded_traffic()
        line = {
            "metric": METRIC if (args.arch, args.ftype) == ("mistral-7b", "Q4_K_M") and n_ctx == 512
            else f"{args.arch} {args.ftype} decode tok/s (bs1, ctx{n_ctx})", "value": value, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8 x q4/q6 integer dots, f32/f64 accumulate",
            "data": "synthetic", "parity_checked": bool(parity), "parity_tokens": parity,
            "config": {"workload": f"{args.arch} {args.ftype} batch-1 greedy decode, {args.prompt}-token prompt + "
                                   f"{args.gen} generated = {n_ctx}-token context (BASELINE.json configs[1])",
                       "tokens_per_step": args.gen, "l2": f"inputs larger than L2 ({info.weight_bytes_per_token / 1e9:.2f} GB weights per token vs 126 MB)",
                       "parallelism": f"replicas x{world} (no collective)", "seed": SEED,
                       "bytes_per_token": bytes_per_token, "timing": "CUDA events on the engine stream, summed over steps",
                       "wall_ms_per_step": wall_ms_max / args.steps},
            "roofline": {"bound": "hbm", "achieved": gemv_gbs, "peak": peak, "unit": "GB/s", "frac": gemv_gbs / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "kernel": "gemv_db_kernel / gemv_db_mixed_kernel (fused norm + act-quant + dequant-GEMV + epilogue)",
                         "method": "in-graph: algorithmic bytes of the GEMV launches of one token / their summed "
                                   "completion-to-completion time from the kernels' own %globaltimer marks (hand-over, "
                                   "dependency wait and activation prologue included)",
                         "launches_per_token": ig["n_gemv"],
                         "algorithmic_bytes_per_token": gemv_bytes_per_token,
                         "avg_launch_us": ig["us_gemv"] / max(1, ig["n_gemv"]),
                         "in_graph_us_per_token": {"gemv": ig["us_gemv"], "attention": ig["us_attn"], "other": ig["us_other"],
                                                   "traced_token": ig["us_token"], "graph_token": us_token},
                         "kernel_time_fits_in_token": bool(fits),
                         "isolated": {"achieved": iso_gbs, "frac": iso_gbs / peak, "avg_launch_us": prof.ms_gemv * 1e3 / max(1, n_gemv),
                                      "attention_ms_per_token": prof.reserved[0] / prof.n_tokens,
                                      "other_ms_per_token": prof.reserved[1] / prof.n_tokens,
                                      "method": "CUDA-event pair around every launch of 16 eagerly launched tokens"},
                         "whole_token_frac_of_peak": bytes_per_token * value / world / 1e9 / peak},
            "e2e": {"value": e2e_value, "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 4 * args.gen,
                    "api": "aios_seq_prefill(last prompt token, from host memory) + aios_seq_step per token "
                           "(one C-ABI call, one 4-byte D2H and one sync per token)"},
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if cpu and world == 1:
            line["cpu_baseline"] = cpu
        if config4:
            line["config4"] = config4
        if config3:
            line["config3"] = config3
        if served:
            line["served"] = served
        print(json.dumps(line), flush=True)
    if eng is not None:
        seq.free()
        seq_e2e.free()
        model.unload()
        eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

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


def cpu_reference_tok_s(path: str, n_prompt: int, max_decode: int, budget_s: float = 15.0, threads: int = 0):
    """Times the CPU restatement of the llama.cpp CPU path (oracle, AVX2 dot kernels) on host cores: greedy decode
    tokens until `budget_s` seconds of CPU work are spent (at least 8, at most `max_decode` tokens)."""
    from oracle import oracle as orc

    orc.build()
    m = orc.OracleModel(path, n_ctx=max(64, n_prompt + max_decode + 1), simd=True, threads=threads)
    cores = orc.lib().oracle_num_threads()
    prompt = gs.synthetic_prompt("mistral-7b", n_prompt, SEED)
    m.eval(prompt)  # prefill (also pages the file in)
    toks = np.array([1], np.int32)
    t0 = time.perf_counter()
    pos, n = n_prompt, 0
    while n < max_decode and (n < 8 or time.perf_counter() - t0 < budget_s):
        logits = m.eval(toks, n_past=pos)
        toks = np.array([int(np.argmax(logits))], np.int32)
        pos += 1
        n += 1
    dt = time.perf_counter() - t0
    m.close()
    return n / dt, cores, dt, n


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    path = ensure_model(args.arch, args.ftype, True)
    from oracle import oracle as orc

    orc.build()
    n_tok = args.ref_tokens_per_step
    m = orc.OracleModel(path, n_ctx=64 + (args.steps + args.warmup) * n_tok + 8, simd=True)
    cores = orc.lib().oracle_num_threads()
    prompt = gs.synthetic_prompt(args.arch, 8, SEED)
    m.eval(prompt)
    pos = len(prompt)
    tok = np.array([1], np.int32)

    def step():
        nonlocal pos, tok
        for _ in range(n_tok):
            lg = m.eval(tok, n_past=pos)
            tok = np.array([int(np.argmax(lg))], np.int32)
            pos += 1

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = args.steps * n_tok / dt
    sample = (f"{args.steps} steps x {n_tok} greedy decode tokens at context {len(prompt)}..{pos} on the same seeded "
              f"{args.arch} {args.ftype} GGUF (bounded sample of the 128-token step)")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int8 x q4/q6 integer dots, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"{args.arch} {args.ftype} batch-1 greedy decode (CPU restatement of the llama.cpp CPU path; "
                               "real llama.cpp is not vendored by the reference and unavailable here)",
                   "tokens_per_step": n_tok},
        "cpu_baseline": {"value": val, "unit": "tok/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
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
    eng = capi.Engine(device=local_rank, max_batch=2)
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

    # roofline leg: per-launch CUDA events around every kernel of a few tokens
    seq.restore()
    prof = seq.profile_decode(16)
    n_gemv = int(prof.reserved[2])
    emb_row = gs.tensor_nbytes((1, info.n_embd), gs.Q4_K if args.ftype == "Q4_K_M" else gs.Q8_0)
    norm_bytes = (2 * info.n_layer + 1) * info.n_embd * 4
    gemv_bytes_per_token = info.weight_bytes_per_token - emb_row  # matrices + the norm vectors their prologues read
    peak, peak_src = measured_peaks()
    gemv_gbs = gemv_bytes_per_token * prof.n_tokens / (prof.ms_gemv * 1e-3) / 1e9
    n_past_mean = args.prompt + (args.gen - 1) / 2.0
    bytes_per_token = info.weight_bytes_per_token + info.kv_bytes_per_position * n_past_mean
    total_tokens = world * args.steps * args.gen
    value = total_tokens / (dev_ms_max * 1e-3)
    e2e_value = total_tokens / (e2e_ms_max * 1e-3)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, cores, dt, n_cpu = cpu_reference_tok_s(path, 8, 512)
            cpu = {"value": v, "unit": "tok/s", "cores": cores, "kind": "port",
                   "sample": f"{n_cpu} greedy decode tokens after an 8-token prompt on the same GGUF ({dt:.1f} s of CPU work); "
                             "CPU restatement with AVX2 integer dots, not llama.cpp (unavailable on this box)"}
        traffic = recorded_traffic()
        line = {
            "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8 x q4/q6 integer dots, f32/f64 accumulate",
            "data": "synthetic",
            "config": {"workload": f"{args.arch} {args.ftype} batch-1 greedy decode, {args.prompt}-token prompt + "
                                   f"{args.gen} generated = {n_ctx}-token context (BASELINE.json configs[1])",
                       "tokens_per_step": args.gen, "l2": "inputs larger than L2 (4.29 GB weights per token vs 126 MB)",
                       "parallelism": f"replicas x{world} (no collective)", "seed": SEED,
                       "bytes_per_token": bytes_per_token, "timing": "CUDA events on the engine stream, summed over steps",
                       "wall_ms_per_step": wall_ms_max / args.steps},
            "roofline": {"bound": "hbm", "achieved": gemv_gbs, "peak": peak, "unit": "GB/s", "frac": gemv_gbs / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "kernel": "gemv_db_kernel / gemv_db_mixed_kernel (fused norm + act-quant + dequant-GEMV + epilogue)", "launches_per_token": n_gemv // prof.n_tokens,
                         "algorithmic_bytes_per_token": gemv_bytes_per_token,
                         "avg_launch_us": prof.ms_gemv * 1e3 / max(1, n_gemv),
                         "whole_token_frac_of_peak": bytes_per_token * value / world / 1e9 / peak,
                         "attention_ms_per_token": prof.reserved[0] / prof.n_tokens,
                         "other_ms_per_token": prof.reserved[1] / prof.n_tokens},
            "e2e": {"value": e2e_value, "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 4 * args.gen,
                    "api": "aios_seq_prefill(last prompt token, from host memory) + aios_seq_step per token "
                           "(one C-ABI call, one 4-byte D2H and one sync per token)"},
            "gpu_launches": launches,
            "clocks": clocks,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    seq.free()
    seq_e2e.free()
    model.unload()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- decode tokens/s of the B200 decode hot path (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--model llama3-70b] [--format nvf4] [--context 8192] [--batch 1] [--configs all|none]

A "step" is one greedy decode step (one token per sequence) of the named model at the named context: embedding
-> all layers (fused RMSNorm, W8A16/NVF4 GEMV, QK-norm + RoPE + KV write, split-KV attention, SwiGLU GEMV,
residual GEMVs) -> final norm -> lm_head GEMV + argmax.

Default workload = BASELINE.json configs[4], north_star's scaling target: Llama-3-70B NVF4 weights, BF16
activations, batch 1, 8k context.  It fits ONE B200 (36 GB of weights), so the same workload is measured at every
N (tensor parallel, strong scaling) and the driver's 1 -> 8 GPU curve measures the claim itself.  The other
single-GPU configurations of BASELINE.json ride in the `configs` object of the same line (each with tokens/s and
the step's HBM roofline fraction): Llama-3-8B FP8 @4k (north_star's >= 70 % target), Qwen3-8B w8a16 @4k
(configs[1]), Llama-3-8B FP8 @32k at 1 / 8 sequences and @16k at 64 sequences (configs[2]; 64 x 32k of KV cache is
275 GB and does not fit one B200), Qwen3-30B-A3B NVF4 (configs[3]).

  value   device-timed tokens/s: K replays of the one-graph step with every input resident in HBM
          (CUDA events on the launching stream, barrier + sync on both sides, max over ranks).
          The weights streamed per step (4.5 .. 41 GB) are far larger than the 126 MB L2, so no flush.
  e2e     the same K steps through the public host-in-the-loop call
          (CausalTransformerModel.decode_step_host): every step copies token/position from pinned
          host memory, replays, and reads the sampled token back.
  roofline   the dominant kernel (gate/up SwiGLU weight-streaming GEMV, in the fused-RMSNorm form the step runs):
          algorithmic bytes / CUDA-event time per launch over the model's own layer weights, against
          MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline   oracle/oracle.c (the C restatement of the reference kernels, OpenMP, all host threads) running
          REAL decoder layers of the same shapes on the host for a bounded time; plus, as `config1_reference`,
          BASELINE.json configs[0] through the reference package itself (oracle/ref_gpt2_cpu.py).
  --impl reference   the CPU arm: K + W real steps, each `layers_per_step` decoder layers + the matching share of
          the lm_head, so that the K timed steps are whole tokens' worth of work; nothing is extrapolated.
N > 1 (torchrun): tensor-parallel over N GPUs (strong scaling: the same token stream, weights sharded).
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

FMT_NAME = {"fp8": "w8a16 FP8-E4M3 128x128 block-scaled", "nvf4": "NVF4 (E2M1, UE4M3 scale per 32)", "bf16": "bf16"}
# (key, model, format, context, batch): the other single-GPU configurations of BASELINE.json
SUB_CONFIGS = [
    ("llama3-8b:fp8:4k", "llama3-8b", "fp8", 4096, 1),
    ("qwen3-8b:fp8:4k", "qwen3-8b", "fp8", 4096, 1),
    ("llama3-8b:fp8:32k", "llama3-8b", "fp8", 32768, 1),
    ("llama3-8b:fp8:32k:b8", "llama3-8b", "fp8", 32768, 8),
    ("llama3-8b:fp8:16k:b64", "llama3-8b", "fp8", 16384, 64),    # 64 x 32k of KV cache (275 GB) exceeds one B200: 16k context
    ("qwen3-30b-a3b:nvf4:4k", "qwen3-30b-a3b", "nvf4", 4096, 1),
]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama3-70b")
    ap.add_argument("--format", default="nvf4", choices=["fp8", "nvf4", "bf16"])
    ap.add_argument("--context", type=int, default=8192)
    ap.add_argument("--batch", type=int, default=1, help="independent sequences decoded per step")
    ap.add_argument("--configs", default="all", help="'all' (default), 'none', or a comma list of sub-config keys")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layers", type=int, default=None, help="debug: truncate the model (INVALID as a bench number)")
    return ap.parse_args()


def measured_peak() -> tuple[float, str]:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ------------------------------------------------------------------ clocks sampling
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int = 0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [ln for (t, ln) in self.lines if t0 - 0.05 <= t <= t1 + 0.15] or [ln for (_, ln) in self.lines]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in rows:
            f = [x.strip() for x in ln.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ CPU legs (oracle port)
class CpuLayer:
    """ONE decoder layer of the named shapes + the lm_head on the host, op by op in the reference's order
    (rmsnorm, q/k/v, [QK-norm], RoPE, KV write, SDPA over the reference's GQA-expanded cache at `context`, o, add,
    rmsnorm, gate, silu, up, mul, down, add; MoE: the top-k experts' SwiGLU), through oracle/oracle.c -- the
    reference has no CPU branch for its FP8 / NVF4 GEMVs (ops/matmul/gemv.py:128, ops/matmul/nvf4.py:191)."""

    def __init__(self, model_name: str, fmt: str, context: int):
        from oracle import oracle as orc
        from pygpukit_b200.llm.config import named_config
        self.orc = orc
        self.cores = orc.set_num_threads(os.cpu_count() or 1)      # torchrun exports OMP_NUM_THREADS=1: override it
        cfg, spec = named_config(model_name, max_seq=context)
        self.cfg, self.spec, self.context = cfg, spec, context
        rng = np.random.default_rng(0)
        H, hd, Hq, Hkv = cfg.hidden_size, cfg.head_dim, cfg.num_heads, cfg.num_kv_heads
        I = cfg.moe_intermediate_size if cfg.is_moe else cfg.intermediate_size

        def rb(shape, lo=0x3800, hi=0x3C00):
            return rng.integers(lo, hi, size=shape, dtype=np.uint16)

        def lin(n_out, n_in):
            if fmt == "fp8":
                codes = rng.integers(0, 0x7E, size=(n_out, n_in), dtype=np.uint8)
                scale = np.full(((n_out + 127) // 128, (n_in + 127) // 128), 0x3A00, np.uint16)
                return lambda x: orc.gemv_fp8_bf16(x, codes, scale)
            if fmt == "nvf4":
                d = rng.integers(0, 256, size=(n_in // 2, n_out), dtype=np.uint8)
                s = np.full((n_in // 32, n_out), 0x30, np.uint8)
                return lambda x: orc.gemv_nvf4_bf16(x, d, s)
            wt = rb((n_out, n_in))
            return lambda x: orc.gemv_bf16(x, wt)

        self.rb = rb
        self.q, self.k, self.v = lin(Hq * hd, H), lin(Hkv * hd, H), lin(Hkv * hd, H)
        self.o, self.gate, self.up, self.down = lin(H, Hq * hd), lin(I, H), lin(I, H), lin(H, I)
        self.n_mlp = cfg.num_experts_per_tok if cfg.is_moe else 1
        self.gamma = np.full(H, 0x3F80, np.uint16)
        self.hgamma = np.full(hd, 0x3F80, np.uint16)
        self.kc, self.vc = rb((Hq, context, hd)), rb((Hq, context, hd))
        self.cos, self.sin = np.ones((1, hd), np.float32), np.zeros((1, hd), np.float32)
        self.hidden = rb((1, H))
        self.head = rb((cfg.vocab_size, H))

    def layer(self):
        orc, cfg = self.orc, self.cfg
        hd, Hq, Hkv = cfg.head_dim, cfg.num_heads, cfg.num_kv_heads
        x = orc.rmsnorm_bf16(self.hidden, self.gamma, cfg.norm_eps)
        qq, kk, vv = self.q(x).reshape(1, Hq, hd), self.k(x).reshape(1, Hkv, hd), self.v(x).reshape(1, Hkv, hd)
        if self.spec.use_qk_norm:
            qq = orc.rmsnorm_bf16(qq.reshape(Hq, hd), self.hgamma, cfg.norm_eps).reshape(1, Hq, hd)
            kk = orc.rmsnorm_bf16(kk.reshape(Hkv, hd), self.hgamma, cfg.norm_eps).reshape(1, Hkv, hd)
        qq, kk = orc.rope_bf16_f32table(qq, kk, self.cos, self.sin)
        orc.kv_cache_update_gqa(kk, self.kc, Hq, self.context - 1)
        orc.kv_cache_update_gqa(vv, self.vc, Hq, self.context - 1)
        att = orc.sdpa_causal_fixed_cache(qq.reshape(Hq, 1, hd), self.kc, self.vc, self.context)
        h = orc.add_inplace_bf16(self.o(att.reshape(1, Hq * hd)), self.hidden)
        x = orc.rmsnorm_bf16(h, self.gamma, cfg.norm_eps)
        m = None
        for _ in range(self.n_mlp):
            g = orc.silu_bf16(self.gate(x))
            m = self.down(orc.mul_inplace_bf16(g, self.up(x)))
        return orc.add_inplace_bf16(m, h)

    def head_rows(self, r0: int, r1: int):
        if r1 > r0:
            lg = self.orc.gemv_bf16(self.hidden, self.head[r0:r1])
            self.orc.argmax_bf16(lg[0])


def cpu_decode_sample(model_name: str, fmt: str, context: int, budget_s: float = 12.0) -> dict:
    """Real decoder layers (+ the lm_head once) on the host for ~budget_s seconds; tokens/s = the measured layer
    rate divided by the model's layer count, with the measured lm_head time added per token."""
    c = CpuLayer(model_name, fmt, context)
    L = c.cfg.num_layers
    c.layer()                                   # warm (page-in, thread pool)
    t0 = time.perf_counter()
    n = 0
    while True:
        c.layer()
        n += 1
        if time.perf_counter() - t0 > budget_s or n >= L:
            break
    t_layer = (time.perf_counter() - t0) / n
    t1 = time.perf_counter()
    c.head_rows(0, c.cfg.vocab_size)
    t_head = time.perf_counter() - t1
    t_tok = L * t_layer + t_head
    return {"value": 1.0 / t_tok, "unit": "tokens/s", "cores": int(c.cores), "kind": "port",
            "sample": f"{n} of the {L} {model_name} {fmt} decoder layers of one token at ctx {context} ({t_layer * 1e3:.1f} ms "
                      f"each, {n * t_layer:.1f} s) + the lm_head once ({t_head * 1e3:.1f} ms), oracle/oracle.c (OpenMP, "
                      f"{c.cores} threads); tokens/s = 1 / ({L} x layer + lm_head)",
            "ms_per_token": t_tok * 1e3}


def gpt2_reference_cpu(prompt_len: int = 16, new_tokens: int = 8) -> dict:
    """BASELINE.json configs[0] through the reference package on its CPU backend (separate process: it sets the
    reference's global backend).  Bounded: the reference's NumPy path runs ~1 token/s and the full 128 + 128 run takes
    minutes, so the default sample is a short prompt + a few greedy tokens, decode timed apart from the prompt pass."""
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref_gpt2_cpu.py"), str(prompt_len),
                            str(new_tokens)], capture_output=True, text=True, timeout=240,
                           env={k: v for k, v in os.environ.items() if k != "OMP_NUM_THREADS"})
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


# ------------------------------------------------------------------ distributed plumbing
def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


class Dist:
    """torch.distributed (NCCL) for barrier / max-over-ranks / unique-id broadcast; a no-op at N == 1."""

    def __init__(self, rank, local_rank, world):
        self.rank, self.local_rank, self.world = rank, local_rank, world
        self.torch = None
        if world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(local_rank)
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
            self.torch, self.dist = torch, dist

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def exchange(self):
        def ex(payload):
            obj = [payload]
            self.dist.broadcast_object_list(obj, src=0)
            return obj[0]
        return ex

    def allgather(self):
        def ag(payload):
            out = [None] * self.world
            self.dist.all_gather_object(out, payload)
            return out
        return ag

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def workload_config(model: str, fmt: str, context: int, batch: int, world: int, layers=None) -> dict:
    return {"workload": f"{model} {FMT_NAME[fmt]} weights, BF16 activations, greedy decode of {batch} sequence(s) at "
                        f"{context} context, random-init weights, synthetic KV cache",
            "shapes": model, "weight_format": fmt, "decode_batch": batch, "context": context,
            "parallelism": f"tp{world}", "l2_hygiene": "per-step weight stream >> 126 MB L2 (no flush needed)",
            **({"layers_truncated_INVALID": layers} if layers else {})}


# ------------------------------------------------------------------ the reference arm (CPU)
def run_reference(args) -> None:
    """--impl reference: the reference's algorithm for this path on the host cores (oracle/oracle.c restates its
    kernels; its FP8 / NVF4 GEMVs have no CPU branch upstream and its native module cannot be built here), all host
    threads.  W + K REAL steps are executed; a step = `lps` decoder layers + lps/L of the lm_head rows, so the timed
    region is K * lps / L whole tokens' worth of work (one layer's synthetic weights are reused: a layer is 0.2-0.5 GB,
    far beyond the CPU caches).  value = tokens' worth of work / elapsed -- measured, not extrapolated."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    t_start = time.perf_counter()
    c = CpuLayer(args.model, args.format, args.context)
    L, V = c.cfg.num_layers, c.cfg.vocab_size
    K, W = max(1, args.steps), max(0, args.warmup)
    c.layer()                                                   # page-in + thread pool (untimed, before warm-up)
    t0 = time.perf_counter()
    c.layer()
    t_probe = time.perf_counter() - t0
    # as many layers per step as fit ~150 s of total work, at most a whole token per step
    lps = int(max(1, min(L, 150.0 / max(t_probe, 1e-4) / (K + W))))
    rows = V * lps // L
    cursor = 0

    def step():
        nonlocal cursor
        for _ in range(lps):
            c.layer()
        c.head_rows(cursor, min(V, cursor + rows))
        cursor = (cursor + rows) % max(1, V - rows + 1)

    for _ in range(W):
        step()
    t0 = time.perf_counter()
    for _ in range(K):
        step()
    dt = time.perf_counter() - t0
    tokens = K * lps / L
    value = tokens / dt
    sample = (f"{K} timed steps x {lps} of {L} {args.model} {args.format} decoder layers (ctx {args.context}) + {rows} of {V} "
              f"lm_head rows per step = {tokens:.3f} tokens' worth of work in {dt:.1f} s, oracle/oracle.c (OpenMP, "
              f"{c.cores} threads)")
    line = {
        "impl": "reference", "metric": "decode_tokens_per_s", "value": value, "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args.model, args.format, args.context, args.batch, world, args.layers),
        "cpu_baseline": {"value": value, "unit": "tokens/s", "cores": int(c.cores), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "tokens_per_step": lps / L, "layers_per_step": lps,
        "gpu_launches": 0, "wall_s": time.perf_counter() - t_start,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ our arm (B200)
def dominant_kernel(model, api, stream_handle, rounds: int = 3):
    """(name, algorithmic bytes per launch, us per launch): the kernel that moves the most bytes per step, in the
    form the step runs it (RMSNorm fused into the launch), timed with CUDA events over the model's OWN per-layer
    weights (L distinct tensors >> L2), back to back under programmatic dependent launch like the step itself."""
    from pygpukit_b200.core.stream import CudaEvent, event_elapsed_ms
    b = model.buffers
    cfg = model.config
    H = cfg.hidden_size
    x = b.hidden if b.hidden.shape[0] == 1 else b.hidden.slice_rows(1)
    act = b.mlp_act if b.mlp_act.shape[0] == 1 else b.mlp_act.slice_rows(1)
    calls, name, nbytes = [], "", 0
    mlp = model.blocks[0].mlp
    norm0 = model.blocks[0].mlp_norm
    if type(mlp).__name__ == "MLP" and getattr(mlp, "gate_up_proj", None) is not None and mlp.fuses_norm(norm0):
        I2 = mlp.gate_up_proj.out_features
        name = "rmsnorm_gemv_nvf4_swiglu (gate/up: fused RMSNorm + SiLU*up)"
        nbytes = mlp.gate_up_proj.weight_bytes + 2 * H + 2 * H + I2
        for blk in model.blocks:
            gu, nm = blk.mlp.gate_up_proj, blk.mlp_norm
            calls.append(lambda gu=gu, nm=nm: api.rmsnorm_gemv_nvf4_swiglu_bf16(
                x.device_ptr, nm.weight.device_ptr, float(nm.eps), gu.data.device_ptr, gu.scale.device_ptr,
                act.device_ptr, 1, H, I2, stream_handle))
    elif type(mlp).__name__ == "MLP" and type(getattr(mlp, "gate_proj", None)).__name__ == "LinearFP8" and mlp.fuses_norm(norm0):
        I = mlp.gate_proj.out_features
        name = "rmsnorm_gemv_fp8_swiglu (gate/up: fused RMSNorm + SiLU*up)"
        nbytes = mlp.gate_proj.weight_bytes + mlp.up_proj.weight_bytes + 2 * H + 2 * H + 2 * I
        for blk in model.blocks:
            g, u, nm = blk.mlp.gate_proj, blk.mlp.up_proj, blk.mlp_norm
            calls.append(lambda g=g, u=u, nm=nm: api.rmsnorm_gemv_fp8_swiglu_bf16(
                x.device_ptr, nm.weight.device_ptr, float(nm.eps), g.weight_fp8.device_ptr, g.scale_inv.device_ptr,
                u.weight_fp8.device_ptr, u.scale_inv.device_ptr, act.device_ptr, 1, H, I, stream_handle))
    else:
        head = model._head
        V = head.shape[0]
        name = "gemv_bf16_argmax (lm_head + greedy)"
        nbytes = head.nbytes + 2 * H + 2 * V
        calls.append(lambda: api.gemv_bf16_argmax(x.device_ptr, head.device_ptr, b.logits.device_ptr,
                                                  b.next_token.device_ptr, H, V, stream_handle))
        rounds = 20
    api.fill_normal_bf16(x.device_ptr, x.size, 4242, 1.0, stream_handle)
    for c in calls:
        c()
    api.stream_sync(stream_handle)
    e0, e1 = CudaEvent(), CudaEvent()
    api.event_record(e0.handle, stream_handle)
    for _ in range(rounds):
        for c in calls:
            c()
    api.event_record(e1.handle, stream_handle)
    e1.synchronize()
    us = event_elapsed_ms(e0, e1) * 1e3 / (rounds * len(calls))
    return name, nbytes, us


def measure(model_name, fmt, ctx, batch, K, W, d, comm, rank, local_rank, peak, *, layers=None, full=False,
            clock_index=None):
    """Build the model, capture the step, time K replays (device-timed) and, when `full`, the host-in-the-loop
    loop and the dominant kernel.  Returns a dict; the model is released before returning."""
    from pygpukit_b200 import llm
    from pygpukit_b200._lib import api, load
    from pygpukit_b200.core.stream import CudaEvent, event_elapsed_ms, use_stream
    lib = load()
    p0 = ctx - (K + W) - 2
    if p0 < 1:
        raise SystemExit(f"--steps + --warmup must stay below --context ({ctx})")
    t_build = time.perf_counter()
    model = llm.build_random_model(model_name, fmt, max_seq=ctx, comm=comm, num_layers=layers)
    if batch > 1:
        model.init_decode(ctx, max_batch_size=1, num_sequences=batch)
    else:
        model.init_decode(ctx, max_batch_size=1)
    for li, blk in enumerate(model.blocks):                       # synthetic KV cache: N(0, 0.1^2)
        for j, c in enumerate((blk.attn._k_cache, blk.attn._v_cache)):
            api.fill_normal_bf16(c.device_ptr, c.size, 7000 + 2 * li + j + 1000 * rank, 0.1, None)
    api.device_sync()
    model.init_decode_graph()
    stream = model._graph_stream
    sh = stream.handle
    t_build = time.perf_counter() - t_build
    toks0 = [(12345 + 977 * s) % model.config.vocab_size for s in range(batch)]

    with use_stream(stream):                                      # kernels per step (one eager step, counted by the library)
        model.set_decode_state(toks0 if batch > 1 else toks0[0], p0)
        lib.pgk_launch_count_reset()
        model.decode_step_device()
        stream.synchronize()
        launches_per_step = int(lib.pgk_launch_count())

    clocks = ClockSampler(clock_index).start() if clock_index is not None else None
    with use_stream(stream):
        model.set_decode_state(toks0 if batch > 1 else toks0[0], p0)
        for _ in range(W):
            model.decode_graph_replay()
        stream.synchronize()
        d.barrier()
        e0, e1 = CudaEvent(), CudaEvent()
        t0 = time.time()
        api.event_record(e0.handle, sh)
        for _ in range(K):
            model.decode_graph_replay()
        api.event_record(e1.handle, sh)
        e1.synchronize()
        t1 = time.time()
        d.barrier()
    ms = d.max(event_elapsed_ms(e0, e1))
    last_tok = [int(t) for t in model.buffers.next_token.to_numpy()[:batch]]
    step_bytes = model.weight_bytes_per_token() + batch * model.kv_bytes_per_token(ctx)
    out = {"tokens_per_s": batch * K / (ms / 1e3), "ms_per_step": ms / K, "launches_per_step": launches_per_step,
           "decode_batch": batch, "context": ctx,
           "step_roofline": {"bytes_per_step_per_gpu": step_bytes, "achieved": step_bytes / (ms / K) / 1e6, "unit": "GB/s",
                             "frac": step_bytes / (ms / K) / 1e6 / peak, "frac_of_8000": step_bytes / (ms / K) / 1e6 / 8000.0},
           "build_s": t_build}
    if full:
        # ---- end to end: host in the loop (pinned H2D of token/position, D2H of the sampled token, every step)
        with use_stream(stream):
            tok, pos = (toks0 if batch > 1 else toks0[0]), p0
            for _ in range(W):
                tok = model.decode_step_host(tok, pos); pos += 1
            d.barrier()
            api.event_record(e0.handle, sh)
            w0 = time.perf_counter()
            for _ in range(K):
                tok = model.decode_step_host(tok, pos); pos += 1
            api.event_record(e1.handle, sh)
            e1.synchronize()
            w1 = time.perf_counter()
            d.barrier()
        e2e_ms = d.max(max(event_elapsed_ms(e0, e1), (w1 - w0) * 1e3))
        got = list(tok) if batch > 1 else [tok]
        nb = 12 * batch
        out["e2e"] = {"value": batch * K / (e2e_ms / 1e3), "unit": "tokens/s", "h2d_bytes_per_step": nb,
                      "d2h_bytes_per_step": 4 * batch, "ms_per_step": e2e_ms / K,
                      "api": "CausalTransformerModel.decode_step_host"}
        out["token_streams_agree"] = bool(got == last_tok)   # both loops start from the same state
        with use_stream(stream):
            kname, kbytes, kus = dominant_kernel(model, api, sh)
        out["dominant"] = (kname, kbytes, kus)
    if clocks is not None:
        out["clocks"] = clocks.stop(t0, t1)
    out["p2p"] = bool(comm is not None and comm._p2p)
    out["fused_exchange"] = bool(getattr(model, "fused_exchange", False))
    # per-op path: do the row-parallel GEMVs of the layers do their own all-reduce (pgk_gemv_*_allreduce_bf16)?
    blk0 = model.blocks[0]
    if not out["fused_exchange"] and getattr(blk0, "comm", None) is not None and blk0.comm.world > 1:
        out["fused_exchange"] = bool(blk0.comm.fuses_allreduce(blk0.attn.o_proj, 1))
    model._graph = None
    api.device_sync()
    del model
    gc.collect()
    api.device_sync()
    return out


def run_ours(args) -> None:
    rank, local_rank, world = dist_env()
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    from pygpukit_b200 import llm
    from pygpukit_b200._lib import api, load
    load()
    api.device_set(local_rank)
    pdl = os.environ.get("PGK_PDL", "1") == "1"
    api.set_pdl(1 if pdl else 0)               # programmatic dependent launch between the step's kernels
    d = Dist(rank, local_rank, world)
    comm = llm.TensorParallel(rank, world).init(exchange=d.exchange(), allgather=d.allgather()) if world > 1 else None
    K, W = args.steps, max(3, args.warmup)
    peak, peak_src = measured_peak()

    main = measure(args.model, args.format, args.context, args.batch, K, W, d, comm, rank, local_rank, peak,
                   layers=args.layers, full=True, clock_index=local_rank if rank == 0 else None)
    kname, kbytes, kus = main["dominant"]
    achieved = kbytes / kus / 1e3
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{args.model}:{args.format}", {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    line = {
        "metric": "decode_tokens_per_s", "value": main["tokens_per_s"], "unit": "tokens/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args.model, args.format, args.context, args.batch, world, args.layers),
        "e2e": main["e2e"],
        "gpu_launches": main["launches_per_step"] * K, "launches_per_step": main["launches_per_step"],
        "roofline": {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "bytes_per_launch": kbytes, "us_per_launch": kus, "frac_of_8000": achieved / 8000.0},
        "step_roofline": main["step_roofline"],
        "pdl": pdl, "allreduce": (("p2p-nvlink" + ("+in-gemv" if main["fused_exchange"] else "")) if main["p2p"]
                                  else ("nccl" if world > 1 else None)),
        "clocks": main.get("clocks"), "token_streams_agree": main["token_streams_agree"], "build_s": main["build_s"],
        "configs": {},
    }

    printed = threading.Event()

    def emit():
        if rank == 0 and not printed.is_set():
            printed.set()
            print(json.dumps(line), flush=True)

    # the other configurations must never cost the headline line: a watchdog prints what there is and leaves
    budget = float(os.environ.get("PGK_BENCH_SUB_BUDGET_S", "240"))
    wd = threading.Timer(budget, lambda: (line["configs"].setdefault("_aborted", f"watchdog after {budget:.0f} s"), emit(),
                                          os._exit(0)))
    wd.daemon = True
    wd.start()
    want = [] if args.configs == "none" else ([s for s in SUB_CONFIGS] if args.configs == "all" else
                                              [s for s in SUB_CONFIGS if s[0] in args.configs.split(",")])
    Ks, Ws = min(K, 32), 4
    for key, mname, fmt, ctx, batch in want:
        if (mname, fmt, ctx, batch) == (args.model, args.format, args.context, args.batch):
            continue
        try:
            r = measure(mname, fmt, ctx, batch, Ks, Ws, d, comm, rank, local_rank, peak)
            line["configs"][key] = {"tokens_per_s": r["tokens_per_s"], "ms_per_step": r["ms_per_step"],
                                    "decode_batch": batch, "context": ctx, "steps": Ks,
                                    "launches_per_step": r["launches_per_step"],
                                    "step_roofline": {"frac": r["step_roofline"]["frac"],
                                                      "achieved": r["step_roofline"]["achieved"], "unit": "GB/s",
                                                      "bytes_per_step_per_gpu": r["step_roofline"]["bytes_per_step_per_gpu"]}}
        except Exception as e:  # noqa: BLE001
            line["configs"][key] = {"error": f"{type(e).__name__}: {e}"[:300]}
            if world > 1:       # a rank that failed alone would leave the others inside a collective: stop here
                break
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        s = cpu_decode_sample(args.model, args.format, args.context, budget_s=12.0)
        line["cpu_baseline"] = {k: s[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line["cpu_baseline"]["config1_reference"] = gpt2_reference_cpu(16, 8)
    wd.cancel()
    emit()
    # teardown: a rank must never hang in a collective destroy after its line is printed
    sys.stdout.flush()
    if world > 1:
        wd2 = threading.Timer(20.0, lambda: os._exit(0))
        wd2.daemon = True
        wd2.start()
    api.device_sync()
    d.barrier()
    if comm is not None:
        comm.destroy()
    d.close()
    if world > 1:
        os._exit(0)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

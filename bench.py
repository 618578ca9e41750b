#!/usr/bin/env python
"""bench.py -- decode tokens/s of the B200 decode hot path (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--model qwen3-8b] [--format fp8] [--context 4096]

A "step" is one greedy decode token (batch 1) of the named model at the named context: embedding
-> all layers (fused RMSNorm, W8A16/NVF4 GEMV, QK-norm + RoPE + KV write, split-KV attention, SwiGLU
GEMV, residual GEMVs) -> final norm -> lm_head GEMV + argmax.  Default workload = BASELINE.json
configs[1]: Qwen3-8B w8a16 (FP8-E4M3 128x128 block-scaled weights, BF16 activations), batch 1,
4k context, random-init weights, synthetic KV cache.

  value   device-timed tokens/s: K replays of the one-graph step with every input resident in HBM
          (CUDA events on the launching stream, barrier + sync on both sides, max over ranks).
          The ~8.8 GB of weights streamed per token are far larger than the 126 MB L2, so no flush.
  e2e     the same K tokens through the public host-in-the-loop call
          (CausalTransformerModel.decode_step_host): every step copies token/position from pinned
          host memory, replays, and reads the sampled token back.
  roofline   the dominant kernel (gate/up SwiGLU weight-streaming GEMV): algorithmic bytes / CUDA-event
          time per launch over the model's own layer weights, against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline   the oracle's C restatement of the reference kernels (oracle/oracle.c, OpenMP) timed on the
          host cores on a bounded sample (one layer + lm_head), extrapolated to tokens/s.
N > 1 (torchrun): tensor-parallel over N GPUs (strong scaling: the same token stream, weights sharded).
"""
from __future__ import annotations

import argparse
import ctypes as C
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="qwen3-8b")
    ap.add_argument("--format", default="fp8", choices=["fp8", "nvf4", "bf16"])
    ap.add_argument("--context", type=int, default=4096)
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


# ------------------------------------------------------------------ CPU baseline (oracle port)
def cpu_decode_sample(model_name: str, fmt: str, context: int, budget_s: float = 15.0) -> dict:
    """Times the oracle's C restatement of the reference kernels (the reference has no CPU branch for
    FP8/NVF4 GEMV: ops/matmul/gemv.py:128, ops/matmul/nvf4.py:191) on ONE decoder layer of the named
    shapes (rmsnorm, q/k/v, [QK-norm], RoPE, KV write, SDPA over the reference's GQA-expanded cache at
    `context`, o, add, rmsnorm, gate, silu, up, mul, down, add) plus the lm_head GEMV + argmax, with
    all host threads, and extrapolates tokens/s = 1 / (L * t_layer + t_head)."""
    from oracle import oracle as orc
    from pygpukit_b200.llm.config import named_config
    cfg, spec = named_config(model_name, max_seq=context)
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
        w = rb((n_out, n_in))
        return lambda x: orc.gemv_bf16(x, w)

    q, k, v = lin(Hq * hd, H), lin(Hkv * hd, H), lin(Hkv * hd, H)
    o, gate, up, down = lin(H, Hq * hd), lin(I, H), lin(I, H), lin(H, I)
    n_mlp = cfg.num_experts_per_tok if cfg.is_moe else 1
    gamma = np.full(H, 0x3F80, np.uint16)
    hgamma = np.full(hd, 0x3F80, np.uint16)
    kc, vc = rb((Hq, context, hd)), rb((Hq, context, hd))
    cos, sin = np.ones((1, hd), np.float32), np.zeros((1, hd), np.float32)
    hidden = rb((1, H))
    pos = context - 1

    def layer():
        x = orc.rmsnorm_bf16(hidden, gamma, cfg.norm_eps)
        qq, kk, vv = q(x).reshape(1, Hq, hd), k(x).reshape(1, Hkv, hd), v(x).reshape(1, Hkv, hd)
        if spec.use_qk_norm:
            qq = orc.rmsnorm_bf16(qq.reshape(Hq, hd), hgamma, cfg.norm_eps).reshape(1, Hq, hd)
            kk = orc.rmsnorm_bf16(kk.reshape(Hkv, hd), hgamma, cfg.norm_eps).reshape(1, Hkv, hd)
        qq, kk = orc.rope_bf16_f32table(qq, kk, cos, sin)
        orc.kv_cache_update_gqa(kk, kc, Hq, pos)
        orc.kv_cache_update_gqa(vv, vc, Hq, pos)
        att = orc.sdpa_causal_fixed_cache(qq.reshape(Hq, 1, hd), kc, vc, context)
        h = orc.add_inplace_bf16(o(att.reshape(1, Hq * hd)), hidden)
        x = orc.rmsnorm_bf16(h, gamma, cfg.norm_eps)
        m = None
        for _ in range(n_mlp):
            g = orc.silu_bf16(gate(x))
            m = down(orc.mul_inplace_bf16(g, up(x)))
        return orc.add_inplace_bf16(m, h)

    layer()                                   # warm (page-in, thread pool)
    t0 = time.perf_counter()
    n = 0
    while True:
        layer()
        n += 1
        if time.perf_counter() - t0 > budget_s * 0.6 or n >= 64:
            break
    t_layer = (time.perf_counter() - t0) / n
    head = rb((cfg.vocab_size, H))
    orc.gemv_bf16(hidden, head[:1024])
    t1 = time.perf_counter()
    lg = orc.gemv_bf16(orc.rmsnorm_bf16(hidden, gamma, cfg.norm_eps), head)
    orc.argmax_bf16(lg[0])
    t_head = time.perf_counter() - t1
    t_tok = cfg.num_layers * t_layer + t_head
    return {"value": 1.0 / t_tok, "unit": "tokens/s", "cores": int(orc.num_threads()), "kind": "port",
            "sample": f"{n} x one {model_name} {fmt} decoder layer at ctx {context} ({t_layer * 1e3:.1f} ms each) + lm_head "
                      f"({t_head * 1e3:.1f} ms), oracle/oracle.c (OpenMP), extrapolated to {cfg.num_layers} layers",
            "ms_per_token": t_tok * 1e3}


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


# ------------------------------------------------------------------ the reference arm (CPU)
def run_reference(args) -> None:
    """--impl reference: the reference's algorithm for this path on the host cores.  Its FP8/NVF4 GEMVs
    have no CPU branch upstream, so this is the oracle's C restatement (kind "port"), all host threads,
    each step a bounded sample (one decoder layer + 1/L of the lm_head time) of the same workload."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    t0 = time.perf_counter()
    budget = min(25.0, max(6.0, 0.2 * (args.steps + args.warmup)))
    s = cpu_decode_sample(args.model, args.format, args.context, budget_s=budget)
    line = {
        "impl": "reference", "metric": "decode_tokens_per_s", "value": s["value"], "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": s["ms_per_token"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {k: s[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": s["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world: int) -> dict:
    fmt_name = {"fp8": "w8a16 FP8-E4M3 128x128 block-scaled", "nvf4": "NVF4 (E2M1, UE4M3 scale per 32)", "bf16": "bf16"}
    return {"workload": f"{args.model} {fmt_name[args.format]} weights, BF16 activations, batch-1 greedy decode at "
                        f"{args.context} context, random-init weights, synthetic KV cache",
            "shapes": args.model, "weight_format": args.format, "decode_batch": 1, "context": args.context,
            "parallelism": f"tp{world}", "l2_hygiene": "per-step weight stream >> 126 MB L2 (no flush needed)",
            **({"layers_truncated_INVALID": args.layers} if args.layers else {})}


# ------------------------------------------------------------------ our arm (B200)
def dominant_kernel(model, api, stream_handle, gk, rounds: int = 3):
    """(name, algorithmic bytes per launch, us per launch): the kernel that moves the most bytes per step,
    timed alone with CUDA events over the model's OWN per-layer weights (L distinct tensors >> L2)."""
    from pygpukit_b200.core.stream import CudaEvent, event_elapsed_ms
    b = model.buffers
    x, cfg = b.norm_out, model.config
    H = cfg.hidden_size
    calls, name, nbytes = [], "", 0
    blk0 = model.blocks[0]
    mlp = blk0.mlp
    if type(mlp).__name__ == "MLP" and mlp.gate_up_proj is not None:
        I2 = mlp.gate_up_proj.out_features
        name = "gemv_nvf4_swiglu (gate/up, fused SiLU*up)"
        nbytes = mlp.gate_up_proj.weight_bytes + 2 * H + I2
        for blk in model.blocks:
            gu = blk.mlp.gate_up_proj
            calls.append(lambda gu=gu: api.gemv_nvf4_swiglu_bf16(x.device_ptr, gu.data.device_ptr, gu.scale.device_ptr,
                                                                 b.mlp_act.device_ptr, 1, H, I2, stream_handle))
    elif type(mlp).__name__ == "MLP" and type(mlp.gate_proj).__name__ == "LinearFP8":
        I = mlp.gate_proj.out_features
        name = "gemv_fp8_swiglu (gate/up, fused SiLU*up)"
        nbytes = mlp.gate_proj.weight_bytes + mlp.up_proj.weight_bytes + 2 * H + 2 * I
        for blk in model.blocks:
            g, u = blk.mlp.gate_proj, blk.mlp.up_proj
            calls.append(lambda g=g, u=u: api.gemv_fp8_swiglu_bf16(
                x.device_ptr, g.weight_fp8.device_ptr, g.scale_inv.device_ptr, u.weight_fp8.device_ptr,
                u.scale_inv.device_ptr, b.mlp_act.device_ptr, 1, H, I, stream_handle))
    else:
        head = model._head
        V = head.shape[0]
        name = "gemv_bf16_argmax (lm_head + greedy)"
        nbytes = head.nbytes + 2 * H + 2 * V
        calls.append(lambda: api.gemv_bf16_argmax(x.device_ptr, head.device_ptr, b.logits.device_ptr,
                                                  b.next_token.device_ptr, H, V, stream_handle))
        rounds = 20
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


def run_ours(args) -> None:
    rank, local_rank, world = dist_env()
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    import pygpukit_b200 as gk
    from pygpukit_b200 import llm
    from pygpukit_b200._lib import api, load
    from pygpukit_b200.core.stream import CudaEvent, event_elapsed_ms, use_stream
    lib = load()
    api.device_set(local_rank)
    pdl = os.environ.get("PGK_PDL", "1") == "1"
    api.set_pdl(1 if pdl else 0)               # programmatic dependent launch between the step's kernels
    d = Dist(rank, local_rank, world)
    comm = llm.TensorParallel(rank, world).init(exchange=d.exchange(), allgather=d.allgather()) if world > 1 else None
    K, W = args.steps, max(3, args.warmup)
    ctx = args.context
    p0 = ctx - (K + W) - 1
    if p0 < 1:
        raise SystemExit(f"--steps + --warmup must stay below --context ({ctx})")

    t_build = time.perf_counter()
    model = llm.build_random_model(args.model, args.format, max_seq=ctx, comm=comm, num_layers=args.layers)
    model.init_decode(ctx, max_batch_size=1)
    for li, blk in enumerate(model.blocks):                       # synthetic KV cache: N(0, 0.1^2)
        for j, c in enumerate((blk.attn._k_cache, blk.attn._v_cache)):
            api.fill_normal_bf16(c.device_ptr, c.size, 7000 + 2 * li + j + 1000 * rank, 0.1, None)
    api.device_sync()
    model.init_decode_graph()
    stream = model._graph_stream
    sh = stream.handle
    t_build = time.perf_counter() - t_build

    # kernels per step (one eager step, counted by the library)
    with use_stream(stream):
        model.set_decode_state(1, p0)
        lib.pgk_launch_count_reset()
        model.decode_step_device()
        stream.synchronize()
        launches_per_step = int(lib.pgk_launch_count())

    # ---- device-timed: K graph replays, inputs resident in HBM
    tok0 = 12345 % model.config.vocab_size
    clocks = ClockSampler(local_rank).start() if rank == 0 else None
    with use_stream(stream):
        model.set_decode_state(tok0, p0)
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
    last_tok = int(model.buffers.next_token.to_numpy()[0])

    # ---- end to end: host in the loop (pinned H2D of token/position, D2H of the sampled token, every step)
    with use_stream(stream):
        tok, pos = tok0, p0
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
    clk = clocks.stop(t0, t1) if clocks else None
    same_tokens = (tok == last_tok)      # both loops start from the same state: the token streams must agree

    # ---- roofline of the dominant kernel + whole-step bytes
    peak, peak_src = measured_peak()
    with use_stream(stream):
        kname, kbytes, kus = dominant_kernel(model, api, sh, gk)
    achieved = kbytes / kus / 1e3
    step_bytes = model.weight_bytes_per_token() + model.kv_bytes_per_token(ctx)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{args.model}:{args.format}", {}).get("dram_bytes_per_launch")
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": "decode_tokens_per_s", "value": K / (ms / 1e3), "unit": "tokens/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args, world),
            "e2e": {"value": K / (e2e_ms / 1e3), "unit": "tokens/s", "h2d_bytes_per_step": 12, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_ms / K, "api": "CausalTransformerModel.decode_step_host"},
            "gpu_launches": launches_per_step * K, "launches_per_step": launches_per_step,
            "roofline": {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_launch": kbytes, "us_per_launch": kus,
                         "frac_of_8000": achieved / 8000.0},
            "step_roofline": {"bytes_per_step_per_gpu": step_bytes, "achieved": step_bytes / (ms / K) / 1e6,
                              "unit": "GB/s", "frac": step_bytes / (ms / K) / 1e6 / peak,
                              "frac_of_8000": step_bytes / (ms / K) / 1e6 / 8000.0},
            "pdl": pdl, "allreduce": ("p2p-nvlink" if (comm is not None and comm._p2p) else ("nccl" if world > 1 else None)),
            "clocks": clk, "token_streams_agree": bool(same_tokens), "build_s": t_build,
        }
        if world == 1 and not args.no_cpu_baseline:
            s = cpu_decode_sample(args.model, args.format, ctx, budget_s=15.0)
            line["cpu_baseline"] = {k: s[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    # teardown: the graph that captured NCCL kernels must go before the communicator does, and a rank must
    # never hang in a collective destroy after its line is printed -- a watchdog ends the process regardless
    sys.stdout.flush()
    if world > 1:
        wd = threading.Timer(20.0, lambda: os._exit(0))
        wd.daemon = True
        wd.start()
    model._graph = None
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
_ = C

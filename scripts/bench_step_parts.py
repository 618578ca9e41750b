"""Per-kernel cost inside the decode step: each op of the step is launched back to back over the
model's OWN per-layer tensors (L distinct weight sets / KV caches, >> L2) and timed with CUDA events.
    python scripts/bench_step_parts.py [model] [format] [context]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.core.stream import CudaEvent, Stream, event_elapsed_ms, use_stream  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "qwen3-8b"
fmt = sys.argv[2] if len(sys.argv) > 2 else "fp8"
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
api.set_pdl(int(os.environ.get("PGK_PDL", "1")))
model = llm.build_random_model(name, fmt, max_seq=ctx)
b = model.init_decode(ctx)
for li, blk in enumerate(model.blocks):
    for j, c in enumerate((blk.attn._k_cache, blk.attn._v_cache)):
        api.fill_normal_bf16(c.device_ptr, c.size, 7000 + 2 * li + j, 0.1, None)
api.device_sync()
st = Stream()
PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0


def timeit(fn, rounds=5):
    with use_stream(st):
        model.set_decode_state(5, ctx - 200)
        for blk in model.blocks:
            fn(blk)
        st.synchronize()
        e0, e1 = CudaEvent(), CudaEvent()
        api.event_record(e0.handle, st.handle)
        for _ in range(rounds):
            for blk in model.blocks:
                fn(blk)
        api.event_record(e1.handle, st.handle)
        e1.synchronize()
    return event_elapsed_ms(e0, e1) * 1e3 / (rounds * len(model.blocks))


H = model.config.hidden_size
hidden = b.hidden
rows = []


def qkv(blk):
    blk.attn.qkv_proj.forward(hidden, out=b.qkv_proj_out, norm=blk.attn_norm)


def qkv_nonorm(blk):
    blk.attn.qkv_proj.forward(hidden, out=b.qkv_proj_out)


def rope(blk):
    a = blk.attn
    api.qk_norm_rope_kv_write(b.qkv_proj_out.device_ptr, None if a.q_norm is None else a.q_norm.weight.device_ptr,
                              None if a.k_norm is None else a.k_norm.weight.device_ptr, 1e-6, a._cos_dev.device_ptr,
                              a._sin_dev.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr, 1, a.num_heads,
                              a.num_kv_heads, a.head_dim, a._max_cache_len, 0, b.position_buf.device_ptr, st.handle)


def sdpa(blk):
    a = blk.attn
    width = (a.num_heads + 2 * a.num_kv_heads) * a.head_dim
    api.sdpa_decode(b.qkv_proj_out.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr, b.attn_flat.device_ptr,
                    a.num_heads, a.num_kv_heads, 1, a.head_dim, a._max_cache_len, ctx, b.context_buf.device_ptr, 0.0,
                    a.head_dim, width, a.head_dim, a.num_heads * a.head_dim, 2, st.handle)


def fused_attn(blk):
    a = blk.attn
    api.rope_sdpa_decode(b.qkv_proj_out.device_ptr, None if a.q_norm is None else a.q_norm.weight.device_ptr,
                         None if a.k_norm is None else a.k_norm.weight.device_ptr, 1e-6, a._cos_dev.device_ptr,
                         a._sin_dev.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr, b.attn_flat.device_ptr,
                         a.num_heads, a.num_kv_heads, a.head_dim, a._max_cache_len, 0, b.position_buf.device_ptr, 0,
                         b.context_buf.device_ptr, 0.0, st.handle)


def oproj(blk):
    blk.attn.o_proj.forward(b.attn_flat, out=b.partial, residual=hidden)


def gateup(blk):
    blk.mlp.gate_up_act(hidden, out=b.mlp_act, norm=blk.mlp_norm)


def down(blk):
    blk.mlp.down_proj.forward(b.mlp_act, out=b.partial, residual=hidden)


def full_layer(blk):
    blk.decode(hidden, 0, ctx, b, b.position_buf, b.context_buf)


a0 = model.blocks[0].attn
kvb = 2 * a0.num_kv_heads * ctx * a0.head_dim * 2
parts = [("qkv (+norm)", qkv, model.blocks[0].attn.qkv_proj.weight_bytes), ("qkv (no norm)", qkv_nonorm, model.blocks[0].attn.qkv_proj.weight_bytes),
         ("qknorm+rope+kv", rope, 0), ("sdpa_decode", sdpa, kvb), ("rope+sdpa fused;", fused_attn, kvb), ("o_proj (+res)", oproj, model.blocks[0].attn.o_proj.weight_bytes)]
if not model.config.is_moe:
    m0 = model.blocks[0].mlp
    gub = m0.gate_up_proj.weight_bytes if m0.gate_up_proj is not None else m0.gate_proj.weight_bytes + m0.up_proj.weight_bytes
    parts += [("gate/up swiglu (+norm)", gateup, gub), ("down (+res)", down, m0.down_proj.weight_bytes)]
def pair_qa(blk):
    qkv(blk); fused_attn(blk)


def pair_ao(blk):
    fused_attn(blk); oproj(blk)


def trio(blk):
    qkv(blk); fused_attn(blk); oproj(blk)


def pair_qo(blk):
    qkv(blk); oproj(blk)


def seq4(blk):
    qkv(blk); fused_attn(blk); oproj(blk); gateup(blk)


def seq5(blk):
    qkv(blk); fused_attn(blk); oproj(blk); gateup(blk); down(blk)


def seq_aogd(blk):
    fused_attn(blk); oproj(blk); gateup(blk); down(blk)


def seq_agd(blk):
    fused_attn(blk); gateup(blk); down(blk)


def seq_ag(blk):
    fused_attn(blk); gateup(blk)


def seq_ad(blk):
    fused_attn(blk); down(blk)


parts += [("seq q>a>o>g", seq4, 0), ("seq q>a>o>g>d", seq5, 0), ("seq a>o>g>d", seq_aogd, 0), ("seq a>g>d", seq_agd, 0), ("seq a>g", seq_ag, 0), ("seq a>d", seq_ad, 0)]
parts += [("seq qkv>attn", pair_qa, 0), ("seq attn>oproj", pair_ao, 0), ("seq qkv>attn>oproj", trio, 0), ("seq qkv>oproj", pair_qo, 0)]
parts.append(("full layer", full_layer, model.blocks[0].weight_bytes + kvb))
if os.environ.get("PARTS"):
    keep = os.environ["PARTS"].split(",")
    parts = [p_ for p_ in parts if any(k in p_[0] for k in keep)]
tot = 0.0
print(f"{name} {fmt} ctx {ctx}  (us per launch, averaged over {len(model.blocks)} layers; peak {PEAK:.0f} GB/s)")
for label, fn, nbytes in parts:
    us = timeit(fn)
    gbs = nbytes / us / 1e3 if nbytes else 0.0
    print(f"  {label:26s} {us:8.2f} us   {nbytes / 1e6:8.2f} MB   {gbs:7.0f} GB/s  {gbs / PEAK:5.2f}")
    if label not in ("full layer", "qkv (no norm)", "qknorm+rope+kv", "sdpa_decode"):
        tot += us
print(f"  sum of parts {tot:.1f} us")

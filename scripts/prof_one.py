"""Launch ONE op of the decode step a few times (for `ncu --set full -k regex:<kernel>`):
    python scripts/prof_one.py sdpa|fattn|qkv|oproj|gateup|down|rope|tc16|tc64 [model] [format] [context]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402

op = sys.argv[1]
name = sys.argv[2] if len(sys.argv) > 2 else "qwen3-8b"
fmt = sys.argv[3] if len(sys.argv) > 3 else "fp8"
ctx = int(sys.argv[4]) if len(sys.argv) > 4 else 4096
model = llm.build_random_model(name, fmt, max_seq=ctx, num_layers=4)
b = model.init_decode(ctx)
for li, blk in enumerate(model.blocks):
    for j, c in enumerate((blk.attn._k_cache, blk.attn._v_cache)):
        api.fill_normal_bf16(c.device_ptr, c.size, 7000 + 2 * li + j, 0.1, None)
model.set_decode_state(5, ctx - 200)
hidden = b.hidden
for rep in range(2):
    for blk in model.blocks:
        a = blk.attn
        if op == "sdpa":
            width = (a.num_heads + 2 * a.num_kv_heads) * a.head_dim
            api.sdpa_decode(b.qkv_proj_out.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr,
                            b.attn_flat.device_ptr, a.num_heads, a.num_kv_heads, 1, a.head_dim, a._max_cache_len, ctx,
                            b.context_buf.device_ptr, 0.0, a.head_dim, width, a.head_dim, a.num_heads * a.head_dim, 2, None)
        elif op == "fattn":
            api.rope_sdpa_decode(b.qkv_proj_out.device_ptr, None if a.q_norm is None else a.q_norm.weight.device_ptr,
                                 None if a.k_norm is None else a.k_norm.weight.device_ptr, 1e-6, a._cos_dev.device_ptr,
                                 a._sin_dev.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr,
                                 b.attn_flat.device_ptr, a.num_heads, a.num_kv_heads, a.head_dim, a._max_cache_len, 0,
                                 b.position_buf.device_ptr, 0, b.context_buf.device_ptr, 0.0, None)
        elif op == "qkv":
            a.qkv_proj.forward(hidden, out=b.qkv_proj_out, norm=blk.attn_norm)
        elif op == "oproj":
            a.o_proj.forward(b.attn_flat, out=b.partial, residual=hidden)
        elif op == "gateup":
            blk.mlp.gate_up_act(hidden, out=b.mlp_act, norm=blk.mlp_norm)
        elif op == "down":
            blk.mlp.down_proj.forward(b.mlp_act, out=b.partial, residual=hidden)
        elif op.startswith("tc"):                   # tc16 / tc64: the tcgen05 skinny GEMM on the gate projection weights
            import pygpukit_b200 as gk
            Mb = int(op[2:])
            g = blk.mlp.gate_proj
            if not hasattr(model, "_tc_x"):
                model._tc_x = gk.empty((Mb, g.in_features), gk.bfloat16)
                api.fill_normal_bf16(model._tc_x.device_ptr, model._tc_x.size, 3, 1.0, None)
                model._tc_y = gk.empty((Mb, g.out_features), gk.bfloat16)
            api.gemm_fp8_bf16_tc(model._tc_x.device_ptr, g.weight_fp8.device_ptr, g.scale_inv.device_ptr,
                                 model._tc_y.device_ptr, Mb, g.in_features, g.out_features, None, None, None)
        elif op == "rope":
            api.qk_norm_rope_kv_write(b.qkv_proj_out.device_ptr, None if a.q_norm is None else a.q_norm.weight.device_ptr,
                                      None if a.k_norm is None else a.k_norm.weight.device_ptr, 1e-6, a._cos_dev.device_ptr,
                                      a._sin_dev.device_ptr, a._k_cache.device_ptr, a._v_cache.device_ptr, 1, a.num_heads,
                                      a.num_kv_heads, a.head_dim, a._max_cache_len, 0, b.position_buf.device_ptr, None)
api.device_sync()
print("ok")

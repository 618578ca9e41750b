"""GPU parity of the one-launch step kernel (csrc/step.cu, pgk_step_*): every GEMV-type op as a ONE-op program
must be bit-identical to the per-op entry point it mirrors (same per-row instruction sequence); the NVF4
row-packed op and the attention op are held to the oracle with the north_star tolerances (2e-3 f32-accumulate
GEMV, 1e-2 BF16 attention); whole models decode the oracle's greedy tokens through it."""
import numpy as np
import pytest

import pygpukit_b200 as gk
from oracle import oracle as orc
from oracle.model import OracleModel
from pygpukit_b200 import llm, ops
from pygpukit_b200._lib import api
from pygpukit_b200.llm import step_program as sp
from pygpukit_b200.llm.layers.linear import LinearBF16, LinearFP8, LinearNVF4
from tests.util import bf16, f32, f32acc_mismatch, rand_bf16, rand_fp8_weights, rel_err

pytestmark = pytest.mark.gpu


def _state(token=0, pos=0, ctx=1):
    st = gk.from_numpy(np.array([pos, ctx, token], np.int32))
    return st, st.narrow(2, 1), st.narrow(0, 1), st.narrow(1, 1), gk.zeros((1,), gk.int32)


def _run(op_list, vocab=1 << 30, token=0, pos=0, ctx=1, advance=False, keep=()):
    st, tok, pb, cb, nxt = _state(token, pos, ctx)
    prog = sp.StepProgram(op_list, tok, pb, cb, nxt, vocab, advance, keep=list(keep) + [st])
    prog.run(None)
    assert prog.error() == 0
    return prog, st, nxt


@pytest.mark.parametrize("K,N", [(4096, 4096), (4096, 6144), (12288, 4096), (14336, 4096), (2048, 520), (8192, 1024),
                                 (1024, 300), (5120, 640), (256, 256)])
@pytest.mark.parametrize("variant", ["plain", "residual", "norm", "norm+residual"])
def test_step_gemv_fp8_bit_identical_to_per_op(K, N, variant):
    rng = np.random.default_rng(K + N)
    codes, scale = rand_fp8_weights(rng, N, K)
    lin = LinearFP8(gk.from_numpy(codes), gk.from_numpy(scale))
    x = gk.from_numpy(rand_bf16(rng, (1, K)))
    res = gk.from_numpy(rand_bf16(rng, (1, N))) if "residual" in variant else None
    gamma = gk.from_numpy(rand_bf16(rng, (K,))) if "norm" in variant else None
    exp = gk.empty((1, N), gk.bfloat16)
    if gamma is not None:
        xn = ops.rmsnorm(x, gamma, 1e-5)                  # two launches == the fused prologue, bit for bit
        api.gemv_fp8_bf16_ex(xn.device_ptr, lin.weight_fp8.device_ptr, lin.scale_inv.device_ptr, exp.device_ptr, 1, K, N,
                             None, None if res is None else res.device_ptr, None)
    else:
        api.gemv_fp8_bf16_ex(x.device_ptr, lin.weight_fp8.device_ptr, lin.scale_inv.device_ptr, exp.device_ptr, 1, K, N,
                             None, None if res is None else res.device_ptr, None)
    out = gk.zeros((1, N), gk.bfloat16)
    _run([sp.gemv_op(lin, x, out, residual=res, gamma=gamma, eps=1e-5)], keep=[lin, x, res, gamma, out])
    assert np.array_equal(out.to_numpy(), exp.to_numpy())


@pytest.mark.parametrize("K,I", [(4096, 12288), (4096, 14336), (8192, 1024), (2048, 768), (1024, 520)])
def test_step_gemv_fp8_swiglu_bit_identical_to_per_op(K, I):
    rng = np.random.default_rng(K + I)
    g = LinearFP8(*[gk.from_numpy(a) for a in rand_fp8_weights(rng, I, K)])
    u = LinearFP8(*[gk.from_numpy(a) for a in rand_fp8_weights(rng, I, K)])
    x = gk.from_numpy(rand_bf16(rng, (1, K)))
    gamma = gk.from_numpy(rand_bf16(rng, (K,)))
    exp = gk.empty((1, I), gk.bfloat16)
    xn = ops.rmsnorm(x, gamma, 1e-6)
    api.gemv_fp8_swiglu_bf16(xn.device_ptr, g.weight_fp8.device_ptr, g.scale_inv.device_ptr, u.weight_fp8.device_ptr,
                             u.scale_inv.device_ptr, exp.device_ptr, 1, K, I, None)
    out = gk.zeros((1, I), gk.bfloat16)
    _run([sp.gemv_op(g, x, out, gamma=gamma, eps=1e-6, flags=sp.F_SWIGLU, lin_up=u)], keep=[g, u, x, gamma, out])
    assert np.array_equal(out.to_numpy(), exp.to_numpy())


@pytest.mark.parametrize("K,V", [(4096, 32000), (2048, 5001), (8192, 3000), (512, 2048)])
def test_step_lm_head_argmax_bit_identical_to_per_op(K, V):
    rng = np.random.default_rng(K + V)
    head = LinearBF16(gk.from_numpy(rand_bf16(rng, (V, K), 0.05)))
    x = gk.from_numpy(rand_bf16(rng, (1, K)))
    gamma = gk.from_numpy(rand_bf16(rng, (K,)))
    exp, exp_tok = gk.empty((1, V), gk.bfloat16), gk.zeros((1,), gk.int32)
    api.rmsnorm_gemv_bf16_argmax(x.device_ptr, gamma.device_ptr, 1e-6, head.weight.device_ptr, exp.device_ptr,
                                 exp_tok.device_ptr, K, V, None) if K >= 512 else None
    out = gk.zeros((1, V), gk.bfloat16)
    _, st, nxt = _run([sp.gemv_op(head, x, out, gamma=gamma, eps=1e-6, flags=sp.F_ARGMAX)], vocab=V, pos=5, ctx=6, advance=True,
                      keep=[head, x, gamma, out])
    assert np.array_equal(out.to_numpy(), exp.to_numpy())
    assert int(nxt.to_numpy()[0]) == int(exp_tok.to_numpy()[0]) == int(np.argmax(f32(out.to_numpy()[0])))
    assert st.to_numpy().tolist() == [6, 7, int(exp_tok.to_numpy()[0])]      # position / context advanced, token fed back


def _nvf4(rng, K, N):
    w = gk.from_numpy(rand_bf16(rng, (K, N), 0.05))
    return LinearNVF4.from_bf16(w)


@pytest.mark.parametrize("K,N", [(8192, 1024), (28672, 256), (3584, 520), (512, 1024), (1024, 2048)])
@pytest.mark.parametrize("residual", [False, True])
def test_step_gemv_nvf4_rows_vs_oracle(K, N, residual):
    rng = np.random.default_rng(K + N)
    lin = _nvf4(rng, K, N)
    x = rand_bf16(rng, (1, K))
    ref = orc.gemv_nvf4_bf16(x, lin.data.to_numpy(), lin.scale.to_numpy(), 1.0)
    res = rand_bf16(rng, (1, N)) if residual else None
    if residual:
        ref = orc.add_inplace_bf16(ref, res)
    packed = sp.pack_nvf4_rows(lin)
    d, s = lin.data.to_numpy(), lin.scale.to_numpy()
    assert np.array_equal(packed.to_numpy(), np.concatenate([d.T, s.T], axis=1))     # the packing is a pure relayout
    out = gk.zeros((1, N), gk.bfloat16)
    gx, gres = gk.from_numpy(x), (gk.from_numpy(res) if residual else None)
    _run([sp.gemv_op(lin, gx, out, residual=gres, packed=packed)], keep=[lin, packed, gx, gres, out])
    assert f32acc_mismatch(out.to_numpy(), ref, 2e-3) == 0
    assert rel_err(out.to_numpy(), ref) < 1e-2


@pytest.mark.parametrize("K,I", [(8192, 1024), (512, 768), (4096, 512)])
def test_step_gemv_nvf4_swiglu_vs_per_op_and_oracle(K, I):
    """gate/up from the interleaved fused tensor (groups of 8 columns) -> two row-packed matrices; result vs the per-op
    fused kernel on the interleaved tensor (same math, different summation order) and vs the oracle chain."""
    rng = np.random.default_rng(K + I)
    gu = _nvf4(rng, K, 2 * I)
    x = rand_bf16(rng, (1, K))
    gamma = rand_bf16(rng, (K,))
    gx, gg = gk.from_numpy(x), gk.from_numpy(gamma)
    pg, pu = sp.pack_nvf4_rows(gu, I, 0, 8, 16), sp.pack_nvf4_rows(gu, I, 8, 8, 16)
    out = gk.zeros((1, I), gk.bfloat16)
    _run([sp.gemv_op(gu, gx, out, gamma=gg, eps=1e-5, flags=sp.F_SWIGLU, packed=pg, packed_up=pu)], keep=[gu, pg, pu, gx, gg, out])
    xn = orc.rmsnorm_bf16(x, gamma, 1e-5)
    full = orc.gemv_nvf4_bf16(xn, gu.data.to_numpy(), gu.scale.to_numpy(), 1.0).reshape(I // 8, 2, 8)
    g, u = full[:, 0, :].reshape(1, I), full[:, 1, :].reshape(1, I)
    ref = orc.mul_inplace_bf16(orc.silu_bf16(g), u)
    # a flipped bf16 rounding of gate or up moves the product by up to 2^-8 relative: BF16 bar
    assert rel_err(out.to_numpy(), ref) < 1e-2
    assert np.mean(out.to_numpy() != ref) < 0.1


def _attn_case(rng, hq, n_kv, max_seq, pos, qk_norm):
    hd = 128
    qkv = rand_bf16(rng, (1, (hq + 2 * n_kv) * hd))
    kc, vc = rand_bf16(rng, (n_kv, max_seq, hd), 0.5), rand_bf16(rng, (n_kv, max_seq, hd), 0.5)
    from pygpukit_b200.llm.layers.rope import precompute_freqs_cis
    cos, sin = precompute_freqs_cis(hd, max_seq, 1e6)
    qg = rand_bf16(rng, (hd,)) if qk_norm else None
    kg = rand_bf16(rng, (hd,)) if qk_norm else None
    return qkv, kc, vc, np.ascontiguousarray(cos, np.float32), np.ascontiguousarray(sin, np.float32), qg, kg


@pytest.mark.parametrize("hq,n_kv", [(32, 8), (64, 8), (8, 1), (4, 1), (8, 2), (2, 2)])
@pytest.mark.parametrize("pos", [0, 6, 15, 16, 99, 1000, 4095])
@pytest.mark.parametrize("qk_norm", [False, True])
def test_step_attention_vs_per_op_kernel(hq, n_kv, pos, qk_norm):
    """The attention op (QK-norm + RoPE + KV write + split-KV attention + fold) against pgk_rope_sdpa_decode on the same
    inputs: same rounding points, different split / fold order -> BF16 outputs within 1e-2 of max (north_star), most
    elements equal; the K/V rows written to the cache are bit-identical."""
    if pos > 1000 and (hq, n_kv) not in ((32, 8), (8, 1)):
        pytest.skip("long contexts on two head layouts")
    rng = np.random.default_rng(hq * 131 + n_kv * 17 + pos)
    max_seq = 4096
    qkv, kc, vc, cos, sin, qg, kg = _attn_case(rng, hq, n_kv, max_seq, pos, qk_norm)
    g = {k: (None if v is None else gk.from_numpy(v)) for k, v in dict(qkv=qkv, cos=cos, sin=sin, qg=qg, kg=kg).items()}
    kc1, vc1, kc2, vc2 = gk.from_numpy(kc), gk.from_numpy(vc), gk.from_numpy(kc), gk.from_numpy(vc)
    exp = gk.zeros((1, hq * 128), gk.bfloat16)
    api.rope_sdpa_decode(g["qkv"].device_ptr, None if qg is None else g["qg"].device_ptr, None if kg is None else g["kg"].device_ptr,
                         1e-6, g["cos"].device_ptr, g["sin"].device_ptr, kc1.device_ptr, vc1.device_ptr, exp.device_ptr,
                         hq, n_kv, 128, max_seq, pos, None, pos + 1, None, 0.0, None)
    out = gk.zeros((1, hq * 128), gk.bfloat16)
    op = sp.StepOp()
    op.kind = sp.OP_ATTN
    op.act, op.out = g["qkv"].device_ptr, out.device_ptr
    op.q_gamma, op.k_gamma = (None if qg is None else g["qg"].device_ptr), (None if kg is None else g["kg"].device_ptr)
    op.attn_eps = 1e-6
    op.cos_table, op.sin_table = g["cos"].device_ptr, g["sin"].device_ptr
    op.k_cache, op.v_cache = kc2.device_ptr, vc2.device_ptr
    op.hq, op.n_kv, op.max_seq, op.scale = hq, n_kv, max_seq, 0.0
    _run([op], pos=pos, ctx=pos + 1, keep=[g, kc2, vc2, out])
    a, b = out.to_numpy(), exp.to_numpy()
    assert np.array_equal(kc2.to_numpy(), kc1.to_numpy()) and np.array_equal(vc2.to_numpy(), vc1.to_numpy())
    assert rel_err(a, b) < 1e-2
    assert np.mean(a != b) < 0.2


def test_step_two_ops_chain_and_repeat():
    """Two dependent ops (grid barrier between them), replayed: the second run sees the first run's state words
    (step parity flips) and gives the same bits."""
    rng = np.random.default_rng(1)
    K, N = 4096, 4096
    l1 = LinearFP8(*[gk.from_numpy(a) for a in rand_fp8_weights(rng, N, K)])
    l2 = LinearFP8(*[gk.from_numpy(a) for a in rand_fp8_weights(rng, K, N)])
    x = gk.from_numpy(rand_bf16(rng, (1, K)))
    mid, out = gk.zeros((1, N), gk.bfloat16), gk.zeros((1, K), gk.bfloat16)
    e1, e2 = gk.empty((1, N), gk.bfloat16), gk.empty((1, K), gk.bfloat16)
    api.gemv_fp8_bf16_ex(x.device_ptr, l1.weight_fp8.device_ptr, l1.scale_inv.device_ptr, e1.device_ptr, 1, K, N, None, None, None)
    api.gemv_fp8_bf16_ex(e1.device_ptr, l2.weight_fp8.device_ptr, l2.scale_inv.device_ptr, e2.device_ptr, 1, N, K, None, None, None)
    prog, st, _ = _run([sp.gemv_op(l1, x, mid), sp.gemv_op(l2, mid, out)], keep=[l1, l2, x, mid, out])
    assert np.array_equal(out.to_numpy(), e2.to_numpy())
    for _ in range(5):
        api.memset(out.device_ptr, 0, out.nbytes, None)
        prog.run(None)
    assert prog.error() == 0
    assert np.array_equal(out.to_numpy(), e2.to_numpy())


MAX_SEQ = 160


@pytest.mark.parametrize("name,fmt", [("small-qwen3", "fp8"), ("small-qwen3", "nvf4")])
def test_step_model_tokens_match_oracle_and_graph_equals_eager(name, fmt):
    model = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=7, std=0.05)
    model.use_step_program = True
    model.init_decode(MAX_SEQ, max_batch_size=16)
    rng = np.random.default_rng(3)
    prompt = [int(t) for t in rng.integers(0, model.config.vocab_size, 12)]
    om = OracleModel(model, MAX_SEQ)
    exp, margins = om.generate(prompt, 128)
    got = model.generate(prompt, 128, temperature=0.0, use_graph=True)
    assert model.step_program() is not None, model.step_unsupported_reason
    assert model.step_program().error() == 0
    for i, (a, e) in enumerate(zip(got[len(prompt):], exp[len(prompt):])):
        if a != e:
            # a divergence is a tie only when the oracle's two best logits are equal or ADJACENT BF16 numbers (one flipped
            # rounding of either swaps them); it ends the comparison
            ulp = 2.0 ** (np.floor(np.log2(max(abs(om.top_logits[i]), 1e-30))) - 7)
            assert margins[i] <= ulp, f"token {i}: got {a}, oracle {e}, margin {margins[i]} (> one BF16 ulp {ulp})"
            assert i >= 32
            break
    model2 = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=7, std=0.05)
    model2.use_step_program = True
    model2.init_decode(MAX_SEQ, max_batch_size=16)
    assert model2.generate(prompt, 128, temperature=0.0, use_graph=False) == got
    # the per-op path decodes the same tokens up to a logit tie
    model3 = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=7, std=0.05)
    model3.use_step_program = False
    model3.init_decode(MAX_SEQ, max_batch_size=16)
    got3 = model3.generate(prompt, 128, temperature=0.0, use_graph=True)
    same = next((i for i, (a, c) in enumerate(zip(got, got3)) if a != c), len(got))
    assert same >= len(prompt) + 32


def test_step_program_rejects_what_it_cannot_run():
    model = llm.build_random_model("tiny-llama", "fp8", max_seq=64, seed=1, std=0.05)      # head_dim 64
    model.use_step_program = True
    model.init_decode(64)
    assert model.step_program() is None and "head_dim" in model.step_unsupported_reason
    model = llm.build_random_model("tiny-moe", "fp8", max_seq=64, seed=1, std=0.05)
    model.use_step_program = True
    model.init_decode(64)
    assert model.step_program() is None

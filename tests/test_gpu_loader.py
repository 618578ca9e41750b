"""Checkpoint loading (pygpukit_b200/llm/loader.py, mirror of llm/loader.py:132-612): models written under the Hugging
Face tensor names are loaded back bit for bit and decode the same tokens; block-wise FP8 checkpoints are detected from
config.json + `weight_scale_inv` tensors; a GPT-2 checkpoint (Conv1D weights, combined c_attn) reproduces the
reference's own CPU run of the same weights (tests/golden/golden_gpt2.npz)."""
import os

import numpy as np
import pytest

from pygpukit_b200 import llm
from pygpukit_b200.core.dtypes import f32_to_bf16_bits
from pygpukit_b200.llm.safetensors import save_safetensors

pytestmark = pytest.mark.gpu
MAX_SEQ = 96


@pytest.mark.parametrize("name,fmt", [("small-qwen3", "fp8"), ("small-qwen3", "bf16"), ("small-llama", "bf16")])
def test_round_trip_same_weights_same_tokens(name, fmt, tmp_path):
    model = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=3, std=0.05)
    path = llm.save_model_to_safetensors(model, str(tmp_path))
    loaded = llm.load_model_from_safetensors(path, max_position_embeddings=MAX_SEQ)
    assert loaded.spec.name == model.spec.name and loaded.weight_format == fmt
    a0, a1 = model.blocks[1].attn, loaded.blocks[1].attn
    if fmt == "fp8":
        assert isinstance(a1.qkv_proj, llm.LinearFP8)
        assert np.array_equal(a0.qkv_proj.weight_fp8.to_numpy(), a1.qkv_proj.weight_fp8.to_numpy())
        assert np.array_equal(a0.qkv_proj.scale_inv.to_numpy(), a1.qkv_proj.scale_inv.to_numpy())
        assert np.array_equal(model.blocks[0].mlp.down_proj.weight_fp8.to_numpy(), loaded.blocks[0].mlp.down_proj.weight_fp8.to_numpy())
    else:
        assert np.array_equal(a0.qkv_proj.weight.to_numpy(), a1.qkv_proj.weight.to_numpy())
    prompt = [3, 141, 59, 26, 535, 89, 79, 323]
    model.init_decode(MAX_SEQ, max_batch_size=8)
    loaded.init_decode(MAX_SEQ, max_batch_size=8)
    assert loaded.generate(prompt, 32, temperature=0.0) == model.generate(prompt, 32, temperature=0.0)


def test_bf16_checkpoint_quantised_at_load(tmp_path):
    """weight_format= re-quantises BF16 linear weights at load time (FP8 blocks / the reference's NVF4 layout)."""
    model = llm.build_random_model("small-qwen3", "bf16", max_seq=MAX_SEQ, seed=3, std=0.05)
    path = llm.save_model_to_safetensors(model, str(tmp_path))
    for fmt, cls in (("fp8", llm.LinearFP8), ("nvf4", llm.LinearNVF4)):
        m = llm.load_model_from_safetensors(path, max_position_embeddings=MAX_SEQ, weight_format=fmt)
        assert isinstance(m.blocks[0].attn.o_proj, cls) and isinstance(m.blocks[0].mlp.down_proj, cls)
        m.init_decode(MAX_SEQ, max_batch_size=8)
        out = m.generate([3, 141, 59, 26], 8, temperature=0.0)
        assert len(out) == 12 and all(0 <= t < m.config.vocab_size for t in out)


def test_gpt2_checkpoint_matches_reference_cpu_run(tmp_path):
    from tests.golden import gpt2_tiny as T
    w, cfg = T.weights(), T.CFG
    H = cfg["n_embd"]

    def b(a):
        return (f32_to_bf16_bits(np.ascontiguousarray(a, np.float32)), "BF16")

    t = {"wte.weight": b(w["wte"]), "wpe.weight": b(w["wpe"]), "ln_f.weight": b(w["ln_f.g"]), "ln_f.bias": b(w["ln_f.b"])}
    for i in range(cfg["n_layer"]):
        p, h = f"h{i}.", f"h.{i}."
        for nm in ("ln_1", "ln_2"):
            t[h + nm + ".weight"], t[h + nm + ".bias"] = b(w[p + nm + ".g"]), b(w[p + nm + ".b"])
        # Hugging Face GPT-2 stores Conv1D weights as [in, out] and q|k|v combined
        t[h + "attn.c_attn.weight"] = b(np.concatenate([w[p + "q.w"], w[p + "k.w"], w[p + "v.w"]], axis=0).T)
        t[h + "attn.c_attn.bias"] = b(np.concatenate([w[p + "q.b"], w[p + "k.b"], w[p + "v.b"]]))
        t[h + "attn.c_proj.weight"], t[h + "attn.c_proj.bias"] = b(w[p + "o.w"].T), b(w[p + "o.b"])
        t[h + "mlp.c_fc.weight"], t[h + "mlp.c_fc.bias"] = b(w[p + "fc1.w"].T), b(w[p + "fc1.b"])
        t[h + "mlp.c_proj.weight"], t[h + "mlp.c_proj.bias"] = b(w[p + "fc2.w"].T), b(w[p + "fc2.b"])
    path = os.path.join(tmp_path, "model.safetensors")
    save_safetensors(path, t)
    model = llm.load_model_from_safetensors(path, config={"model_type": "gpt2", **cfg})
    assert model.spec.name == "gpt2" and model.config.hidden_size == H
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_gpt2.npz"))
    toks = [int(x) for x in g["tokens"]]
    model.init_decode(cfg["n_positions"], max_batch_size=16)
    got = model.generate(list(T.PROMPT), T.N_NEW, temperature=0.0)
    agreed = 0
    for i in range(T.N_NEW):
        if got[len(T.PROMPT) + i] != toks[len(T.PROMPT) + i]:
            assert g["margins"][i] <= 0.12, f"token {i} differs at a clear margin {g['margins'][i]}"
            break
        agreed += 1
    assert agreed >= 12

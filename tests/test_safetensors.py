"""safetensors reader / writer (pygpukit_b200/llm/safetensors.py): CPU-only round trips, including BF16 / FP8 bit
patterns and a sharded index (the reference's reader is its Rust loader; behaviour per llm/safetensors.py)."""
import json
import os

import numpy as np
import pytest

from pygpukit_b200.llm.safetensors import load_safetensors, save_safetensors


def test_round_trip_single_file(tmp_path):
    rng = np.random.default_rng(0)
    t = {
        "a.weight": (rng.integers(0, 65536, (5, 8), dtype=np.uint16), "BF16"),
        "b.codes": (rng.integers(0, 255, (4, 16), dtype=np.uint8), "F8_E4M3"),
        "c": (rng.standard_normal((3, 2, 2)).astype(np.float32), "F32"),
        "d": (rng.standard_normal((7,)).astype(np.float16), "F16"),
        "e": (np.arange(6, dtype=np.int32).reshape(2, 3), "I32"),
    }
    path = os.path.join(tmp_path, "model.safetensors")
    save_safetensors(path, t, {"format": "pt"})
    st = load_safetensors(path)
    assert sorted(st.tensor_names) == sorted(t)
    for name, (arr, dt) in t.items():
        info = st.tensor_info(name)
        assert info.dtype == dt and info.shape == arr.shape
        assert np.array_equal(st.tensor(name), arr)
    with open(path, "rb") as f:                          # header is what the format says it is
        n = int.from_bytes(f.read(8), "little")
        header = json.loads(f.read(n))
    assert header["__metadata__"] == {"format": "pt"} and header["c"]["dtype"] == "F32"


def test_sharded_index(tmp_path):
    a = (np.arange(12, dtype=np.uint16).reshape(3, 4), "BF16")
    b = (np.arange(8, dtype=np.uint8).reshape(2, 4), "U8")
    save_safetensors(os.path.join(tmp_path, "model-00001-of-00002.safetensors"), {"x": a})
    save_safetensors(os.path.join(tmp_path, "model-00002-of-00002.safetensors"), {"y": b})
    idx = os.path.join(tmp_path, "model.safetensors.index.json")
    with open(idx, "w") as f:
        json.dump({"weight_map": {"x": "model-00001-of-00002.safetensors", "y": "model-00002-of-00002.safetensors"}}, f)
    st = load_safetensors(idx)
    assert np.array_equal(st.tensor("x"), a[0]) and np.array_equal(st.tensor("y"), b[0])
    assert "x" in st and "z" not in st


def test_rejects_unknown_dtype(tmp_path):
    path = os.path.join(tmp_path, "bad.safetensors")
    hdr = json.dumps({"t": {"dtype": "C64", "shape": [1], "data_offsets": [0, 8]}}).encode()
    with open(path, "wb") as f:
        f.write(len(hdr).to_bytes(8, "little") + hdr + b"\0" * 8)
    with pytest.raises(ValueError):
        load_safetensors(path)


def test_llm_exports_legacy_names():
    """Names of pygpukit/llm/__init__.py that are aliases upstream (models/causal.py:1496-1501) resolve to the same classes."""
    from pygpukit_b200 import llm
    assert llm.RMSNorm is llm.Norm and llm.LayerNorm is llm.Norm
    assert llm.LlamaAttention is llm.Attention and llm.CausalSelfAttention is llm.Attention
    assert llm.LlamaMLP is llm.MLP and llm.LlamaBlock is llm.TransformerBlock
    assert llm.ShardedSafeTensorsFile is llm.SafeTensorsFile
    for name in ("load_mixtral_from_safetensors", "TensorInfo", "DecodeSpeculative", "DecodeJacobi", "sample_token"):
        assert hasattr(llm, name)

"""safetensors reader / writer in pure Python + NumPy (mirror of pygpukit/llm/safetensors.py; the reference's
reader is its Rust mmap loader rust/pygpukit-core/src/llm/tensor_loader.rs, which is out of scope -- SURVEY.md 8f-3).

File format: 8-byte little-endian header length, a JSON header {name: {dtype, shape, data_offsets}}, then the raw
tensor bytes.  Tensors are memory-mapped; `tensor(name)` returns a NumPy view (BF16 and FP8 as uint16 / uint8 bit
patterns, which is what the device arrays of this package hold).  A `*.index.json` path opens every shard it lists.
"""
from __future__ import annotations

import json
import os
import struct
from dataclasses import dataclass

import numpy as np

# safetensors dtype -> (numpy storage dtype, bytes per element)
_DTYPES = {
    "F64": (np.float64, 8), "F32": (np.float32, 4), "F16": (np.float16, 2), "BF16": (np.uint16, 2),
    "F8_E4M3": (np.uint8, 1), "F8_E5M2": (np.uint8, 1), "I64": (np.int64, 8), "I32": (np.int32, 4),
    "I16": (np.int16, 2), "I8": (np.int8, 1), "U8": (np.uint8, 1), "BOOL": (np.bool_, 1),
}


@dataclass(frozen=True)
class TensorInfo:
    name: str
    dtype: str
    shape: tuple[int, ...]
    offset: int
    nbytes: int
    shard: int


class SafeTensorsFile:
    """One file or a sharded checkpoint (`model.safetensors.index.json`)."""

    def __init__(self, path: str):
        self.path = path
        self._maps: list[np.memmap] = []
        self._info: dict[str, TensorInfo] = {}
        files = [path]
        if path.endswith(".index.json"):
            with open(path, encoding="utf-8") as f:
                index = json.load(f)
            base = os.path.dirname(os.path.abspath(path))
            files = [os.path.join(base, fn) for fn in sorted(set(index["weight_map"].values()))]
        for shard, fn in enumerate(files):
            with open(fn, "rb") as f:
                (hlen,) = struct.unpack("<Q", f.read(8))
                header = json.loads(f.read(hlen).decode("utf-8"))
            data0 = 8 + hlen
            self._maps.append(np.memmap(fn, dtype=np.uint8, mode="r"))
            for name, meta in header.items():
                if name == "__metadata__":
                    continue
                if meta["dtype"] not in _DTYPES:
                    raise ValueError(f"{fn}: tensor {name!r} has unsupported dtype {meta['dtype']}")
                b0, b1 = meta["data_offsets"]
                self._info[name] = TensorInfo(name, meta["dtype"], tuple(int(d) for d in meta["shape"]), data0 + b0, b1 - b0, shard)

    @property
    def tensor_names(self) -> list[str]:
        return list(self._info)

    def __contains__(self, name: str) -> bool:
        return name in self._info

    def tensor_info(self, name: str) -> TensorInfo:
        return self._info[name]

    def tensor(self, name: str) -> np.ndarray:
        """Zero-copy view of the tensor's bytes in its storage dtype (BF16 -> uint16 bits, FP8 -> uint8 codes)."""
        t = self._info[name]
        npdt, esz = _DTYPES[t.dtype]
        n = int(np.prod(t.shape)) if t.shape else 1
        if n * esz != t.nbytes:
            raise ValueError(f"tensor {name!r}: {t.nbytes} bytes for shape {t.shape} {t.dtype}")
        raw = self._maps[t.shard][t.offset:t.offset + t.nbytes]
        return raw.view(npdt).reshape(t.shape)


def load_safetensors(path: str) -> SafeTensorsFile:
    return SafeTensorsFile(path)


def save_safetensors(path: str, tensors: dict[str, tuple[np.ndarray, str]], metadata: dict[str, str] | None = None) -> None:
    """tensors: name -> (array, safetensors dtype string); BF16 / FP8 arrays are their bit patterns."""
    header: dict = {}
    if metadata:
        header["__metadata__"] = metadata
    off = 0
    blobs = []
    for name, (arr, dt) in tensors.items():
        npdt, esz = _DTYPES[dt]
        a = np.ascontiguousarray(arr).view(npdt) if arr.dtype != npdt else np.ascontiguousarray(arr)
        nbytes = a.size * esz
        header[name] = {"dtype": dt, "shape": list(arr.shape), "data_offsets": [off, off + nbytes]}
        off += nbytes
        blobs.append(a)
    hbytes = json.dumps(header, separators=(",", ":")).encode("utf-8")
    hbytes += b" " * ((8 - len(hbytes) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hbytes)))
        f.write(hbytes)
        for a in blobs:
            f.write(a.tobytes())


__all__ = ["SafeTensorsFile", "TensorInfo", "load_safetensors", "save_safetensors"]

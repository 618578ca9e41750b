"""Streams, events and CUDA-graph capture over the C ABI.

Replaces pygpukit/core/stream.py and the pybind CudaGraph / CudaEvent classes
(native/bindings/core_bindings.cpp:409-509).  Every op in pygpukit_b200.ops is enqueued on
the *current* stream (default: the legacy default stream) and is asynchronous;
GPUArray.to_numpy() synchronises.  (The reference syncs after every op,
native/ops/common/error.cuh:28-37 -- behaviourally identical for a single-threaded caller.)
"""
from __future__ import annotations

import contextlib
import ctypes as C

from .._lib import api

_current = None  # ctypes void* handle or None (default stream)


def current_stream_handle():
    return _current


class Stream:
    def __init__(self) -> None:
        h = C.c_void_p()
        api.stream_create(C.byref(h))
        self.handle = h

    def synchronize(self) -> None:
        api.stream_sync(self.handle)

    def __enter__(self) -> "Stream":
        global _current
        self._prev = _current
        _current = self.handle
        return self

    def __exit__(self, *exc) -> None:
        global _current
        _current = self._prev

    def __del__(self) -> None:
        try:
            api.stream_destroy(self.handle)
        except Exception:
            pass


@contextlib.contextmanager
def use_stream(stream: "Stream | None"):
    global _current
    prev = _current
    _current = stream.handle if stream is not None else None
    try:
        yield
    finally:
        _current = prev


def default_stream():
    return None


def synchronize() -> None:
    """Wait for the current stream (device-wide when it is the default stream)."""
    if _current is None:
        api.device_sync()
    else:
        api.stream_sync(_current)


class CudaEvent:
    def __init__(self) -> None:
        h = C.c_void_p()
        api.event_create(C.byref(h))
        self.handle = h

    def record(self, stream: "Stream | None" = None) -> None:
        api.event_record(self.handle, stream.handle if stream is not None else _current)

    def synchronize(self) -> None:
        api.event_sync(self.handle)

    def __del__(self) -> None:
        try:
            api.event_destroy(self.handle)
        except Exception:
            pass


def event_elapsed_ms(start: CudaEvent, stop: CudaEvent) -> float:
    ms = C.c_float()
    api.event_elapsed_ms(start.handle, stop.handle, C.byref(ms))
    return float(ms.value)


def event_elapsed_us(start: CudaEvent, stop: CudaEvent) -> float:
    return event_elapsed_ms(start, stop) * 1000.0


class CudaGraph:
    """Capture everything enqueued inside `with graph.capture(stream):` and replay it.

    Mirrors the reference's CudaGraph (begin_capture / end_capture / replay,
    llm/decode/m1_graph.py:248-457) on one explicit stream.
    """

    def __init__(self) -> None:
        self._exec = None
        self._stream: Stream | None = None

    @contextlib.contextmanager
    def capture(self, stream: Stream):
        self._stream = stream
        with use_stream(stream):
            api.graph_begin(stream.handle)
            try:
                yield self
            finally:
                h = C.c_void_p()
                api.graph_end(stream.handle, C.byref(h))
                self._exec = h

    def replay(self, stream: "Stream | None" = None) -> None:
        if self._exec is None:
            raise RuntimeError("CudaGraph.replay: nothing captured")
        s = stream if stream is not None else self._stream
        api.graph_launch(self._exec, s.handle if s is not None else None)

    launch = replay

    def __del__(self) -> None:
        try:
            if self._exec is not None:
                api.graph_destroy(self._exec)
        except Exception:
            pass

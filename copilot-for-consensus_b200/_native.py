"""ctypes binding of the C-ABI library (include/b200_retrieval.h).

There is deliberately no fallback: if ``libb200_retrieval.so`` has not been built, or no sm_100
device is visible, importing the native layer / creating a handle fails loudly.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200_retrieval.so")

OK, ERR_INVALID, ERR_DIMENSION, ERR_NOT_FOUND, ERR_CUDA, ERR_UNSUPPORTED, ERR_NOMEM = 0, -1, -2, -3, -4, -5, -6
MASTER_F32, MASTER_BF16 = 0, 1


class EncoderConfigC(ctypes.Structure):
    _fields_ = [
        ("vocab_size", ctypes.c_int32), ("hidden", ctypes.c_int32), ("layers", ctypes.c_int32),
        ("heads", ctypes.c_int32), ("intermediate", ctypes.c_int32), ("max_pos", ctypes.c_int32),
        ("type_vocab", ctypes.c_int32), ("ln_eps", ctypes.c_float),
    ]


#: every symbol include/b200_retrieval.h declares: name -> (restype, argtypes)
_vp, _i, _i64, _cp = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_char_p
SYMBOLS = {
    "b200_last_error": (_cp, []),
    "b200_version": (_i, []),
    "b200_device_count": (_i, []),
    "b200_set_option": (_i, [_cp, _i]),
    "b200_get_option": (_i, [_cp, ctypes.POINTER(_i)]),
    "b200_index_create": (_i, [_i, _i64, _i, _i, ctypes.POINTER(_vp)]),
    "b200_index_destroy": (_i, [_vp]),
    "b200_index_clear": (_i, [_vp]),
    "b200_index_count": (_i64, [_vp]),
    "b200_index_dim": (_i, [_vp]),
    "b200_index_reserve": (_i, [_vp, _i64]),
    "b200_index_add": (_i, [_vp, _vp, _i64, _i, _vp]),
    "b200_index_remove_rows": (_i, [_vp, _vp, _i64, _vp]),
    "b200_index_get_rows": (_i, [_vp, _vp, _i64, _vp, _i, _vp]),
    "b200_index_search": (_i, [_vp, _vp, _i, _i, _i64, _i, _i64, _vp, _vp, _i, _vp]),
    "b200_index_certificate_stats": (_i, [_vp, ctypes.POINTER(_i64), ctypes.POINTER(_i64)]),
    "b200_index_search_l2": (_i, [_vp, _vp, _i, _i, _i64, _i, _i64, _vp, _vp, _i, _vp]),
    "b200_index_search_static": (_i, [_vp, _vp, _i, _i64, _i, _i64, _vp, _vp, _vp]),
    "b200_index_search_packed": (_i, [_vp, _vp, _i, _i64, _i, _i64, _vp, _vp]),
    "b200_merge_hits": (_i, [_vp, _i, _i64, _i, _i64, _i64, _i, _vp, _vp, _i, _vp]),
    "b200_index_search_exchange": (_i, [_vp, _vp, _i, _i64, _i, _i64, _vp, _vp, _i, _i, ctypes.c_uint32, _vp]),
    "b200_merge_hits_wait": (_i, [_vp, _vp, ctypes.c_uint32, _i, _i64, _i, _i64, _i64, _i, _vp, _vp, _i, _vp]),
    "b200_merge_topk": (_i, [_vp, _vp, _i64, _i, _i, _i, _vp, _vp, _i, _vp]),
    "b200_encoder_create": (_i, [ctypes.POINTER(EncoderConfigC), _i, ctypes.POINTER(_vp)]),
    "b200_encoder_destroy": (_i, [_vp]),
    "b200_encoder_load_tensor": (_i, [_vp, _cp, _vp, _i64]),
    "b200_encoder_finalize": (_i, [_vp]),
    "b200_encoder_hidden": (_i, [_vp]),
    "b200_encode": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp]),
    "b200_encoder_peek": (_i, [_vp, _i, _vp, _i64]),
    "b200_wordpiece_create": (_i, [_vp, _vp, ctypes.c_int32, _i, ctypes.POINTER(_vp)]),
    "b200_wordpiece_destroy": (_i, [_vp]),
    "b200_wordpiece_encode": (_i, [_vp, _vp, _vp, _i64, ctypes.c_int32, _vp, _vp, ctypes.c_int32]),
    "b200_profile_enable": (_i, [_i]),
    "b200_profile_reset": (_i, []),
    "b200_profile_read": (_i64, [_vp, _i64]),
}

_lib = None


def lib() -> ctypes.CDLL:
    """Load the shared library once; raise if it is missing (no CPU path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C copilot-for-consensus_b200/csrc`). There is no CPU fallback."
            )
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError here means header and library disagree
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def last_error() -> str:
    msg = lib().b200_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int) -> None:
    """Map a C status to the Python exception the reference's contract uses."""
    if rc == OK:
        return
    msg = last_error() or f"b200 error {rc}"
    if rc in (ERR_INVALID, ERR_DIMENSION):
        raise ValueError(msg)
    if rc == ERR_NOT_FOUND:
        raise KeyError(msg)
    if rc == ERR_NOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)


def device_index(device) -> int:
    """'cuda', 'cuda:3', 3, torch.device -> ordinal.  Anything that is not a CUDA device is refused."""
    if isinstance(device, int):
        return device
    s = str(device)
    if s == "cuda":
        try:
            import torch

            return torch.cuda.current_device() if torch.cuda.is_available() else 0
        except Exception:
            return 0
    if s.startswith("cuda:"):
        return int(s.split(":", 1)[1])
    raise ValueError(f"the B200 path needs a CUDA device, got {device!r}")


def current_stream_ptr(device_ordinal: int) -> int:
    """torch's current stream on that device as a cudaStream_t value (0 = legacy default stream)."""
    try:
        import torch

        if torch.cuda.is_available():
            return int(torch.cuda.current_stream(device_ordinal).cuda_stream)
    except Exception:
        pass
    return 0


_PINNED_MAX_BYTES = 64 << 20


def host_out(shape, dtype, device=None):
    """A numpy array for a result the library copies back from the GPU.  Page-locked when torch can provide it (its
    caching host allocator recycles the blocks of arrays that have been dropped, so after the first calls this costs
    microseconds): the device-to-host copy then runs at the DMA rate instead of being staged through the driver's bounce
    buffer, and so does the host-to-device copy when the caller hands the same array back (embeddings -> query).  A
    fresh array per call, like ``np.empty``; plain pageable memory for results over 64 MB or without torch.  ``device``:
    the ordinal the caller works on -- the allocation is made with that device current, so that a process driving
    ``cuda:3`` does not create a context on ``cuda:0`` for it."""
    import numpy as np

    nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
    if 0 < nbytes <= _PINNED_MAX_BYTES:
        try:
            import torch

            if torch.cuda.is_available():
                tdt = {"float32": torch.float32, "int64": torch.int64, "int32": torch.int32}.get(np.dtype(dtype).name)
                if tdt is not None:
                    if device is None:
                        return torch.empty(tuple(int(x) for x in shape), dtype=tdt, pin_memory=True).numpy()
                    with torch.cuda.device(int(device)):
                        return torch.empty(tuple(int(x) for x in shape), dtype=tdt, pin_memory=True).numpy()
        except Exception:
            pass
    return np.empty(shape, dtype=dtype)


def set_option(name: str, value: int) -> None:
    """Process-wide kernel-variant switch (``b200_set_option``); see the header for names and defaults."""
    check(lib().b200_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    v = ctypes.c_int(0)
    check(lib().b200_get_option(name.encode(), ctypes.byref(v)))
    return int(v.value)


class options:
    """``with options(fuse_ffn=0, gemm_ast=0): ...`` -- set switches for a block and restore them afterwards."""

    def __init__(self, **kw: int):
        self._kw, self._old = kw, {}

    def __enter__(self):
        for k, v in self._kw.items():
            self._old[k] = get_option(k)
            set_option(k, v)
        return self

    def __exit__(self, *exc):
        for k, v in self._old.items():
            set_option(k, v)
        return False


def profile(on: bool) -> None:
    lib().b200_profile_reset()
    lib().b200_profile_enable(1 if on else 0)


def profile_read() -> dict[str, dict[str, float]]:
    """{kernel name: {"launches", "ms", "flops", "bytes"}} accumulated since the last reset."""
    buf = ctypes.create_string_buffer(1 << 16)
    n = lib().b200_profile_read(buf, len(buf))
    out: dict[str, dict[str, float]] = {}
    for line in buf.value[: max(n, 0)].decode().splitlines():
        name, cnt, ms, fl, by = line.split()
        out[name] = {"launches": int(cnt), "ms": float(ms), "flops": float(fl), "bytes": float(by)}
    return out

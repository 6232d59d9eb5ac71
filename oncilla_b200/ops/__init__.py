"""Direct access to the sm_100a kernels for torch tensors / raw device pointers.

Each function enqueues one of OUR kernels (include/ocm_kernels.h) on the
current torch CUDA stream.  There is no eager/PyTorch fallback: on a machine
without a GPU these raise.  ``oracle_*`` are the plain C++ CPU restatements used
as test oracles (src/oracle/).
"""
from __future__ import annotations

import ctypes as C

from .. import _native as N
from .._native import OcmError, lib


def _ptr(t) -> int:
    return int(t) if isinstance(t, int) else int(t.data_ptr())


def _nbytes(t) -> int:
    return t.numel() * t.element_size()


def _stream(stream=None) -> C.c_void_p:
    if stream is None:
        import torch

        stream = torch.cuda.current_stream()
    return C.c_void_p(stream if isinstance(stream, int) else int(stream.cuda_stream))


def _check(r: int, what: str) -> None:
    if r:
        raise OcmError(what, r)


def copy(dst, src, nbytes: int | None = None, *, dst2=None, signal=None, stream=None, mode: int | None = None,
         ctas: int = 0, agents: int = 0, stages: int = 0, tile_bytes: int = 0) -> None:
    """K1/K2/K3: dst[:n] = src[:n] (and dst2[:n]) - local, peer-mapped or pinned-host pointers."""
    n = _nbytes(src) if nbytes is None else int(nbytes)
    d2 = C.c_void_p(_ptr(dst2)) if dst2 is not None else C.c_void_p(0)
    sg = C.c_void_p(_ptr(signal)) if signal is not None else C.c_void_p(0)
    if mode is None and not (ctas or agents or stages or tile_bytes):
        _check(lib.ocmk_copy(C.c_void_p(_ptr(dst)), d2, C.c_void_p(_ptr(src)), n, sg, _stream(stream)), "ocmk_copy")
    else:
        _check(lib.ocmk_copy_tuned(C.c_void_p(_ptr(dst)), d2, C.c_void_p(_ptr(src)), n, sg, _stream(stream),
                                   int(mode or 0), ctas, agents, stages, tile_bytes), "ocmk_copy_tuned")


def fill(dst, pattern: int = 0, nbytes: int | None = None, *, stream=None) -> None:
    n = _nbytes(dst) if nbytes is None else int(nbytes)
    _check(lib.ocmk_fill(C.c_void_p(_ptr(dst)), n, C.c_uint64(pattern & (2**64 - 1)), _stream(stream)), "ocmk_fill")


def verify(t, pattern: int, nbytes: int | None = None) -> int:
    """Number of bytes that differ from the repeating 8-byte pattern."""
    import torch

    out = torch.zeros(1, dtype=torch.int64, device=t.device if hasattr(t, "device") else "cuda")
    n = _nbytes(t) if nbytes is None else int(nbytes)
    _check(lib.ocmk_verify(C.c_void_p(_ptr(t)), n, C.c_uint64(pattern & (2**64 - 1)), C.c_void_p(out.data_ptr()),
                           _stream()), "ocmk_verify")
    return int(out.item())


def checksum(t, nbytes: int | None = None) -> int:
    import torch

    out = torch.zeros(1, dtype=torch.int64, device=t.device if hasattr(t, "device") else "cuda")
    n = _nbytes(t) if nbytes is None else int(nbytes)
    _check(lib.ocmk_checksum(C.c_void_p(_ptr(t)), n, C.c_void_p(out.data_ptr()), _stream()), "ocmk_checksum")
    return int(out.item()) & (2**64 - 1)


def signal(sig, *, stream=None) -> None:
    _check(lib.ocmk_signal(C.c_void_p(_ptr(sig)), _stream(stream)), "ocmk_signal")


def _desc(dst, src, row_bytes, counts, src_strides, dst_strides, src_dtype, dst_dtype, scale_block, scales):
    d = N.ocmk_strided()
    d.src, d.dst, d.row_bytes = _ptr(src), _ptr(dst), int(row_bytes)
    for i in range(3):
        d.n[i] = int(counts[i]) if i < len(counts) else 1
        d.src_stride[i] = int(src_strides[i]) if i < len(src_strides) else 0
        d.dst_stride[i] = int(dst_strides[i]) if i < len(dst_strides) else 0
    d.src_dtype, d.dst_dtype, d.scale_block = int(src_dtype), int(dst_dtype), int(scale_block)
    d.scales = _ptr(scales) if scales is not None else None
    return d


def strided_copy(dst, src, *, row_bytes: int, counts=(1, 1, 1), src_strides=(0, 0, 0), dst_strides=(0, 0, 0),
                 src_dtype: int = N.DT_U8, dst_dtype: int = N.DT_U8, scale_block: int = 0, scales=None, signal=None,
                 stream=None) -> None:
    """K4/K5: gather rows at src strides, scatter at dst strides, optionally casting on the way."""
    d = _desc(dst, src, row_bytes, counts, src_strides, dst_strides, src_dtype, dst_dtype, scale_block, scales)
    sg = C.c_void_p(_ptr(signal)) if signal is not None else C.c_void_p(0)
    _check(lib.ocmk_strided_copy(C.byref(d), sg, _stream(stream)), "ocmk_strided_copy")


def oracle_strided_copy(dst, src, *, row_bytes: int, counts=(1, 1, 1), src_strides=(0, 0, 0), dst_strides=(0, 0, 0),
                        src_dtype: int = N.DT_U8, dst_dtype: int = N.DT_U8, scale_block: int = 0, scales=None) -> None:
    """CPU restatement on host tensors / pointers."""
    d = _desc(dst, src, row_bytes, counts, src_strides, dst_strides, src_dtype, dst_dtype, scale_block, scales)
    if lib.ocm_oracle_strided_copy(C.byref(d)) != 0:
        raise ValueError("oracle rejected the descriptor")


def put_replicated(dsts, src, nbytes: int | None = None, *, stream=None) -> None:
    """K6: one source to up to 8 destinations through a thread-block cluster sharing each staged tile."""
    n = _nbytes(src) if nbytes is None else int(nbytes)
    arr = (C.c_void_p * len(dsts))(*[_ptr(d) for d in dsts])
    _check(lib.ocmk_put_replicated(arr, len(dsts), C.c_void_p(_ptr(src)), n, _stream(stream)), "ocmk_put_replicated")


def gups(bases, words_per_stripe: int, total_words: int, n_updates: int, seed: int = 1, *, stream=None) -> None:
    """K7: n_updates random 8-byte xor updates over a table striped across len(bases) slabs."""
    arr = (C.c_void_p * len(bases))(*[_ptr(b) for b in bases])
    _check(lib.ocmk_gups(arr, len(bases), int(words_per_stripe), int(total_words), int(n_updates), int(seed),
                         _stream(stream)), "ocmk_gups")


def gups_threads(n_updates: int) -> int:
    return int(lib.ocmk_gups_threads(int(n_updates)))


def sort_u64(keys, tmp=None, *, stream=None) -> None:
    """K8: in-place radix sort of an int64/uint64 tensor interpreted as unsigned keys."""
    import torch

    if tmp is None:
        tmp = torch.empty_like(keys)
    _check(lib.ocmk_sort_u64(C.c_void_p(_ptr(keys)), C.c_void_p(_ptr(tmp)), keys.numel(), _stream(stream)),
           "ocmk_sort_u64")


def merge_runs(keys, tmp, offsets, *, stream=None):
    """K8 merge: ``keys`` holds ``len(offsets) - 1`` ascending runs laid end to end (run i = keys[offsets[i]:offsets[i+1]]);
    returns the tensor (``keys`` or ``tmp``) that holds the merged sequence.  Unsigned order, stable."""
    offs = (C.c_uint64 * len(offsets))(*[int(o) for o in offsets])
    r = lib.ocmk_merge_runs_u64(C.c_void_p(_ptr(keys)), C.c_void_p(_ptr(tmp)), offs, len(offsets) - 1, _stream(stream))
    _check(min(r, 0), "ocmk_merge_runs_u64")
    return tmp if r == 1 else keys


def launch_count() -> int:
    return int(lib.ocmk_launch_count())


def _device_guarded(fn):
    """Launch on the device of the first CUDA tensor argument (not on whatever device is current)."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kw):
        dev = None
        for a in list(args) + list(kw.values()):
            if isinstance(a, (list, tuple)) and a:
                a = a[0]
            d = getattr(a, "device", None)
            if d is not None and getattr(d, "type", None) == "cuda":
                dev = d
                break
        if dev is None:
            return fn(*args, **kw)
        import torch

        with torch.cuda.device(dev):
            return fn(*args, **kw)

    return wrapper


for _n in ("copy", "fill", "verify", "checksum", "signal", "strided_copy", "put_replicated", "gups", "sort_u64"):
    globals()[_n] = _device_guarded(globals()[_n])

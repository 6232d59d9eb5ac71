"""Pythonic face of libocm: ``init/tini``, ``Allocation``, copies, signals, stats.

PyTorch is used only to view allocation memory as tensors and to supply
streams; every byte is moved by the native library (reference API surface:
reference inc/oncillamem.h:69-89, implementation src/lib.c).
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

from . import _native as N
from ._native import OcmError, lib

_inited = False


def init() -> None:
    global _inited
    if lib.ocm_init() != 0:
        raise OcmError("ocm_init")
    _inited = True


def tini() -> None:
    global _inited
    if _inited:
        lib.ocm_tini()
        _inited = False


def rank() -> int:
    return lib.ocm_rank()


def num_ranks() -> int:
    return lib.ocm_num_ranks()


def device() -> int:
    return lib.ocm_device()


def prefetch_views(gpu: bool = True, host: bool = True) -> int:
    """Map every rank's pools into this process now (set-up time) instead of at first use; returns how many ranks
    are fully mapped.  Mapping a 100 GiB peer pool takes seconds, once per process."""
    r = lib.ocm_prefetch_views((1 if gpu else 0) | (2 if host else 0))
    if r < 0:
        raise OcmError("ocm_prefetch_views", -r)
    return r


def multicast_ready(wait_s: float = 0.0) -> bool:
    """True once the NVLS multicast region is bound on every GPU and mapped here (daemons bring it up in ~100 ms)."""
    import time

    t0 = time.time()
    while True:
        if lib.ocm_multicast_ready():
            return True
        if time.time() - t0 >= wait_s:
            return False
        time.sleep(0.02)


def quiet() -> None:
    if lib.ocm_quiet() != 0:
        raise OcmError("ocm_quiet", N.lib.ocm_last_error())


def _stream_ptr(stream) -> Optional[int]:
    """None -> library stream; a torch.cuda.Stream or raw int -> its cudaStream_t."""
    if stream is None:
        return None
    h = stream if isinstance(stream, int) else int(stream.cuda_stream)
    # torch's default stream has handle 0, which the C API reads as "library stream":
    # name it explicitly (cudaStreamLegacy == 1)
    return h if h else 1


def current_stream_ptr() -> int:
    import torch

    return int(torch.cuda.current_stream().cuda_stream)


class _CudaView:
    """Minimal __cuda_array_interface__ carrier so torch can alias device memory we own."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}
        self._owner = owner


class Allocation:
    """One OCM allocation (a local buffer, or a pair: local window + far side)."""

    def __init__(self, kind: int, local_bytes: int = 0, remote_bytes: int = 0, *, policy: int = N.POLICY_DEFAULT,
                 rank: int = -1, flags: int = 0):
        p = N.ocm_alloc_params_ex()
        p.base.local_alloc_bytes = int(local_bytes)
        p.base.rem_alloc_bytes = int(remote_bytes)
        p.base.kind = int(kind)
        p.policy = int(policy)
        p.rank = int(rank)
        p.flags = int(flags)
        self._h = lib.ocm_alloc_ex(C.byref(p))
        if not self._h:
            raise OcmError(f"ocm_alloc(kind={N.KIND_NAMES.get(kind, kind)}, local={local_bytes}, remote={remote_bytes})")
        self.requested_kind = kind

    # ---- introspection --------------------------------------------------------------
    @property
    def handle(self):
        if not self._h:
            raise OcmError("allocation already freed", 8)
        return self._h

    @property
    def kind(self) -> int:
        return lib.ocm_alloc_kind(self.handle)

    @property
    def is_remote(self) -> bool:
        return bool(lib.ocm_is_remote(self.handle))

    @property
    def is_multicast(self) -> bool:
        return bool(lib.ocm_is_multicast(self.handle))

    @property
    def remote_rank(self) -> int:
        return lib.ocm_remote_rank(self.handle)

    @property
    def remote_bytes(self) -> int:
        n = C.c_size_t(0)
        if lib.ocm_remote_sz(self.handle, C.byref(n)) != 0:
            return -1
        return n.value

    @property
    def id(self) -> int:
        return lib.ocm_alloc_id(self.handle)

    def localbuf(self) -> tuple[int, int]:
        buf, n = C.c_void_p(0), C.c_size_t(0)
        if lib.ocm_localbuf(self.handle, C.byref(buf), C.byref(n)) != 0:
            return 0, 0
        return int(buf.value or 0), n.value

    @property
    def local_ptr(self) -> int:
        return self.localbuf()[0]

    @property
    def local_bytes(self) -> int:
        return self.localbuf()[1]

    @property
    def remote_ptr(self) -> int:
        return int(lib.ocm_remote_ptr(self.handle) or 0)

    def _is_gpu_local(self) -> bool:
        return self.kind in (N.OCM_LOCAL_GPU, N.OCM_REMOTE_GPU, N.OCM_STRIPED_GPU, N.OCM_REPLICATED_GPU)

    # ---- tensor views ------------------------------------------------------------------
    def local_tensor(self, dtype=None):
        """The local side as a torch tensor aliasing OCM memory (uint8 unless dtype is given)."""
        import torch

        ptr, n = self.localbuf()
        if not ptr:
            raise OcmError("allocation has no local side", 1)
        if self._is_gpu_local():
            t = torch.as_tensor(_CudaView(ptr, n, self), device=f"cuda:{device()}")
        else:
            buf = (C.c_uint8 * n).from_address(ptr)
            t = torch.frombuffer(buf, dtype=torch.uint8)
            t._ocm_owner = self
        return t if dtype is None else t.view(dtype)

    def remote_tensor(self, dtype=None):
        """The far side as mapped into this process (peer HBM: usable by kernels on our GPU)."""
        import torch

        ptr, n = self.remote_ptr, self.remote_bytes
        if not ptr:
            raise OcmError("allocation has no directly mapped far side", 1)
        if self.kind in (N.OCM_REMOTE_GPU,):
            t = torch.as_tensor(_CudaView(ptr, n, self), device=f"cuda:{device()}")
        else:
            buf = (C.c_uint8 * n).from_address(ptr)
            t = torch.frombuffer(buf, dtype=torch.uint8)
            t._ocm_owner = self
        return t if dtype is None else t.view(dtype)

    def slab_tensor(self, rank: int, dtype=None):
        """Multi-slab kinds: the slab that lives on ``rank`` (peer HBM mapped into this process)."""
        import torch

        ptr, n = int(lib.ocm_slab_ptr(self.handle, int(rank)) or 0), int(lib.ocm_slab_bytes(self.handle))
        if not ptr:
            raise OcmError("allocation has no slab on that rank", 1)
        t = torch.as_tensor(_CudaView(ptr, n, self), device=f"cuda:{device()}")
        return t if dtype is None else t.view(dtype)

    # ---- signals -------------------------------------------------------------------------
    def signal(self) -> int:
        v = C.c_uint64(0)
        r = lib.ocm_signal_read(self.handle, C.byref(v))
        if r:
            raise OcmError("ocm_signal_read", r)
        return v.value

    def wait_signal(self, at_least: int, timeout_ms: int = 10000) -> None:
        r = lib.ocm_signal_wait(self.handle, int(at_least), int(timeout_ms))
        if r:
            raise OcmError("ocm_signal_wait", r)

    # ---- sharing between applications ---------------------------------------------------------
    def share(self) -> bytes:
        """A 32-byte token another application of the cluster passes to :func:`attach`."""
        tok = N.ocm_share_token()
        r = lib.ocm_share(self.handle, C.byref(tok))
        if r:
            raise OcmError("ocm_share", r)
        return bytes(tok)

    @property
    def attach_count(self) -> int:
        return lib.ocm_attach_count(self.handle)

    # ---- copies -----------------------------------------------------------------------------
    def put(self, nbytes: int | None = None, *, src_offset: int = 0, dest_offset: int = 0, stream="sync") -> None:
        """One-sided write window[src_offset] -> far side[dest_offset]."""
        _onesided(self, 1, nbytes, src_offset, dest_offset, stream)

    def get(self, nbytes: int | None = None, *, src_offset: int = 0, dest_offset: int = 0, stream="sync") -> None:
        """One-sided read far side[dest_offset] -> window[src_offset]."""
        _onesided(self, 0, nbytes, src_offset, dest_offset, stream)

    def batch(self, entries, *, stream=None) -> None:
        """Many one-sided transfers in ONE kernel launch.  ``entries``: iterable of ``(op, nbytes, src_offset,
        dest_offset)`` with op "put"/1 or "get"/0.  Stream-ordered (``None`` = library stream, wait with :func:`quiet`)."""
        entries = list(entries)
        arr = (N.ocm_params * len(entries))()
        for p, (op, nbytes, so, do) in zip(arr, entries):
            p.op_flag = 1 if op in (1, "put") else 0
            p.bytes, p.src_offset, p.dest_offset = int(nbytes), int(so), int(do)
        r = lib.ocm_copy_onesided_batch(self.handle, arr, len(entries), C.c_void_p(_stream_ptr(stream)))
        if r:
            raise OcmError("ocm_copy_onesided_batch", r)

    def reduce_get_f32(self, nbytes: int | None = None, *, src_offset: int = 0, dest_offset: int = 0, stream=None) -> None:
        """NVLS gather-reduce: window[src_offset:] = fp32 sum over every rank's replica of far[dest_offset:].

        Multicast-backed ``OCM_REPLICATED_GPU`` allocations only; stream-ordered (``None`` = library stream,
        wait with :func:`quiet`)."""
        p = N.ocm_params()
        p.src_offset, p.dest_offset = int(src_offset), int(dest_offset)
        p.bytes = int(self.local_bytes if nbytes is None else nbytes)
        r = lib.ocm_reduce_get_f32(self.handle, C.byref(p), C.c_void_p(_stream_ptr(stream)))
        if r:
            raise OcmError("ocm_reduce_get_f32", r)

    def copy_in(self, src_ptr: int) -> None:
        if lib.ocm_copy_in(self.handle, C.c_void_p(src_ptr)) != 0:
            raise OcmError("ocm_copy_in")

    def copy_out(self, dst_ptr: int) -> None:
        if lib.ocm_copy_out(C.c_void_p(dst_ptr), self.handle) != 0:
            raise OcmError("ocm_copy_out")

    def free(self) -> None:
        if self._h:
            h, self._h = self._h, None
            if lib.ocm_free(h) != 0:
                raise OcmError("ocm_free")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.free()

    def __repr__(self):
        if not self._h:
            return "<Allocation freed>"
        return (f"<Allocation {N.KIND_NAMES.get(self.kind)} local={self.local_bytes} remote={self.remote_bytes} "
                f"rank={self.remote_rank}>")


def alloc(kind: int, local_bytes: int = 0, remote_bytes: int = 0, **kw) -> Allocation:
    return Allocation(kind, local_bytes, remote_bytes, **kw)


def attach(token: bytes) -> Allocation:
    """Open a block another application shared (``Allocation.share()``): a window-less pair whose far side
    is that block.  ``free()`` on it only drops the handle."""
    tok = N.ocm_share_token.from_buffer_copy(token)
    h = C.c_void_p(0)
    r = lib.ocm_attach(C.byref(tok), C.byref(h))
    if r or not h.value:
        raise OcmError("ocm_attach", r)
    a = Allocation.__new__(Allocation)
    a._h = h.value
    a.requested_kind = lib.ocm_alloc_kind(a._h)
    return a


_tls = threading.local()


def _scratch_params():
    """One ``ocm_params`` per thread, reused by the one-sided calls (the library reads it during the call only);
    building a ctypes structure per message costs as much as the library call itself for small messages."""
    try:
        return _tls.p, _tls.ref
    except AttributeError:
        _tls.p = N.ocm_params()
        _tls.ref = C.byref(_tls.p)
        return _tls.p, _tls.ref


def _onesided(a: Allocation, op: int, nbytes, src_offset, dest_offset, stream) -> None:
    p, ref = _scratch_params()
    p.src_offset = int(src_offset)
    p.dest_offset = int(dest_offset)
    p.bytes = int(a.local_bytes if nbytes is None else nbytes)
    p.op_flag = op
    h = a._h
    if not h:
        raise OcmError("allocation already freed", 8)
    if stream == "sync":
        if lib.ocm_copy_onesided(h, ref) != 0:
            raise OcmError("ocm_copy_onesided")
    else:
        r = lib.ocm_copy_onesided_async(h, ref, _stream_ptr(stream))
        if r:
            raise OcmError("ocm_copy_onesided_async", r)


def copy(dst: Allocation, src: Allocation, nbytes: int, *, src_offset: int = 0, dest_offset: int = 0,
         src_offset_2: int = 0, dest_offset_2: int = 0, op_flag: int = 1, stream="sync") -> None:
    """Two-sided ``ocm_copy`` (see DESIGN.md for the offset semantics of each hop)."""
    p = N.ocm_params()
    p.src_offset, p.dest_offset = int(src_offset), int(dest_offset)
    p.src_offset_2, p.dest_offset_2 = int(src_offset_2), int(dest_offset_2)
    p.bytes = int(nbytes)
    p.op_flag = int(op_flag)
    if stream == "sync":
        if lib.ocm_copy(dst.handle, src.handle, C.byref(p)) != 0:
            raise OcmError("ocm_copy")
    else:
        r = lib.ocm_copy_async(dst.handle, src.handle, C.byref(p), C.c_void_p(_stream_ptr(stream)))
        if r:
            raise OcmError("ocm_copy_async", r)


def copy_strided(dst: Allocation, src: Allocation, *, row_bytes: int, counts=(1, 1, 1), src_strides=(0, 0, 0),
                 dst_strides=(0, 0, 0), src_offset: int = 0, dst_offset: int = 0, src_dtype: int = N.DT_U8,
                 dst_dtype: int = N.DT_U8, scale_block: int = 0, scale_offset: int = 0, op_flag: int = 1,
                 onesided: bool = False, stream=None) -> None:
    d = N.ocm_strided()
    d.src_offset, d.dst_offset = int(src_offset), int(dst_offset)
    d.extent[0] = int(row_bytes)
    for i in range(3):
        d.extent[i + 1] = int(counts[i]) if i < len(counts) else 1
        d.src_stride[i] = int(src_strides[i]) if i < len(src_strides) else 0
        d.dst_stride[i] = int(dst_strides[i]) if i < len(dst_strides) else 0
    d.op_flag, d.src_dtype, d.dst_dtype = int(op_flag), int(src_dtype), int(dst_dtype)
    d.scale_block, d.scale_offset = int(scale_block), int(scale_offset)
    sp = C.c_void_p(_stream_ptr(stream))
    if onesided:
        r = lib.ocm_copy_onesided_strided(dst.handle, C.byref(d), sp)
    else:
        r = lib.ocm_copy_strided(dst.handle, src.handle, C.byref(d), sp)
    if r:
        raise OcmError("ocm_copy_strided", r)


def server_stats() -> dict:
    """The resident copy server of this process: enabled?, copies served, kernel (re)starts."""
    ops, launches = C.c_uint64(0), C.c_uint64(0)
    on = lib.ocm_server_stats(C.byref(ops), C.byref(launches))
    return {"enabled": bool(on), "ops": ops.value, "launches": launches.value}


def stats(rank: int | None = None) -> dict:
    s = N.ocm_stats()
    r = lib.ocm_get_stats(C.byref(s)) if rank is None else lib.ocm_get_rank_stats(int(rank), C.byref(s))
    if r:
        raise OcmError("ocm_get_stats", r)
    return {n: getattr(s, n) for n, _ in s._fields_}

"""Locate, build and load the native core (libocm.so, oncillamem) and declare its C ABI.

The native tree is built in place under ``build/`` by CMake (``build_native``);
the shared object is loaded with ctypes.  There is no Python fallback for any
device operation: if the library is missing the import of this module fails
loudly (``OCM_AUTOBUILD=1`` builds it on demand).
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
BUILD = REPO / "build"
LIB_PATH = BUILD / "lib" / "libocm.so"
DAEMON_PATH = BUILD / "bin" / "oncillamem"


def build_native(verbose: bool = False, jobs: int | None = None) -> None:
    """Configure + build every native target for sm_100a (nvcc cross-compiles without a GPU)."""
    cmake = shutil.which("cmake")
    ninja = shutil.which("ninja")
    if cmake is None:
        raise RuntimeError("cmake not found")
    BUILD.mkdir(exist_ok=True)
    env = dict(os.environ)
    env.setdefault("CUDACXX", "/usr/local/cuda/bin/nvcc")
    env["PATH"] = "/usr/local/cuda/bin:" + env.get("PATH", "")
    cache = BUILD / "CMakeCache.txt"
    # a cache configured for another source path (the tree travels between machines) is useless
    if cache.exists() and f"CMAKE_HOME_DIRECTORY:INTERNAL={REPO}" not in cache.read_text(errors="ignore"):
        shutil.rmtree(BUILD)
        BUILD.mkdir()
    gen = ["-G", "Ninja"] if ninja else []
    out = None if verbose else subprocess.PIPE
    r = subprocess.run([cmake, "-S", str(REPO), "-B", str(BUILD), "-DCMAKE_BUILD_TYPE=Release", *gen],
                       env=env, stdout=out, stderr=subprocess.STDOUT, text=True)
    if r.returncode:
        raise RuntimeError("cmake configure failed:\n" + (r.stdout or ""))
    cmd = [cmake, "--build", str(BUILD)]
    if jobs:
        cmd += ["-j", str(jobs)]
    r = subprocess.run(cmd, env=env, stdout=out, stderr=subprocess.STDOUT, text=True)
    if r.returncode:
        raise RuntimeError("native build failed:\n" + (r.stdout or "")[-8000:])


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        if os.environ.get("OCM_AUTOBUILD", "1") == "1":
            build_native()
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} not built - run `python -c 'import __graft_entry__ as g; g.build()'`")
    return C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)


lib = _load()

# ---- types -----------------------------------------------------------------------
ocm_alloc_t = C.c_void_p

OCM_LOCAL_HOST, OCM_LOCAL_RMA, OCM_REMOTE_RMA, OCM_LOCAL_RDMA, OCM_REMOTE_RDMA, OCM_LOCAL_GPU, OCM_REMOTE_GPU, \
    OCM_REMOTE_HOST, OCM_STRIPED_GPU, OCM_REPLICATED_GPU = range(1, 11)
KIND_NAMES = {1: "LOCAL_HOST", 2: "LOCAL_RMA", 3: "REMOTE_RMA", 4: "LOCAL_RDMA", 5: "REMOTE_RDMA", 6: "LOCAL_GPU",
              7: "REMOTE_GPU", 8: "REMOTE_HOST", 9: "STRIPED_GPU", 10: "REPLICATED_GPU"}

POLICY_DEFAULT, POLICY_NEIGHBOUR, POLICY_EXPLICIT, POLICY_LEAST_LOADED, POLICY_ROUND_ROBIN, POLICY_SELF = range(6)
ALLOC_SLOW_PATH, ALLOC_NO_ZERO, ALLOC_NO_WINDOW = 1, 2, 4
DT_U8, DT_F32, DT_BF16, DT_F8E4M3 = 0, 1, 2, 3


class ocm_params(C.Structure):
    _fields_ = [("src_offset", C.c_uint64), ("dest_offset", C.c_uint64), ("src_offset_2", C.c_uint64),
                ("dest_offset_2", C.c_uint64), ("bytes", C.c_uint64), ("op_flag", C.c_int)]


class ocm_alloc_params(C.Structure):
    _fields_ = [("local_alloc_bytes", C.c_uint64), ("rem_alloc_bytes", C.c_uint64), ("kind", C.c_int)]


class ocm_alloc_params_ex(C.Structure):
    _fields_ = [("base", ocm_alloc_params), ("policy", C.c_int), ("rank", C.c_int), ("flags", C.c_uint32),
                ("reserved", C.c_uint32)]


class ocm_strided(C.Structure):
    _fields_ = [("src_offset", C.c_uint64), ("dst_offset", C.c_uint64), ("extent", C.c_uint64 * 4),
                ("src_stride", C.c_int64 * 3), ("dst_stride", C.c_int64 * 3), ("op_flag", C.c_int),
                ("src_dtype", C.c_int), ("dst_dtype", C.c_int), ("scale_block", C.c_int),
                ("scale_offset", C.c_uint64)]


class ocm_share_token(C.Structure):
    _fields_ = [("id", C.c_uint64), ("bytes", C.c_uint64), ("rank", C.c_int32), ("region", C.c_uint32),
                ("generation", C.c_uint32), ("tier", C.c_uint32)]


class ocm_stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "allocs_live", "allocs_total", "frees_total", "bytes_gpu_live", "bytes_host_live", "copies", "copy_bytes",
        "kernel_launches", "fast_path_allocs", "slow_path_allocs", "pool_gpu_bytes", "pool_gpu_used",
        "pool_host_bytes", "pool_host_used", "pool_gpu_granule_bytes", "pool_host_granule_bytes", "table_faults")]


class ocmk_strided(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_uint64), ("n", C.c_uint64 * 3),
                ("src_stride", C.c_int64 * 3), ("dst_stride", C.c_int64 * 3), ("src_dtype", C.c_int),
                ("dst_dtype", C.c_int), ("scale_block", C.c_int), ("scales", C.c_void_p)]


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


P = C.POINTER
_sig("ocm_init", C.c_int)
_sig("ocm_tini", C.c_int)
_sig("ocm_alloc", ocm_alloc_t, P(ocm_alloc_params))
_sig("ocm_alloc_ex", ocm_alloc_t, P(ocm_alloc_params_ex))
_sig("ocm_free", C.c_int, ocm_alloc_t)
_sig("ocm_localbuf", C.c_int, ocm_alloc_t, P(C.c_void_p), P(C.c_size_t))
_sig("ocm_is_remote", C.c_bool, ocm_alloc_t)
_sig("ocm_alloc_kind", C.c_int, ocm_alloc_t)
_sig("ocm_remote_sz", C.c_int, ocm_alloc_t, P(C.c_size_t))
_sig("ocm_copy_out", C.c_int, C.c_void_p, ocm_alloc_t)
_sig("ocm_copy_in", C.c_int, ocm_alloc_t, C.c_void_p)
_sig("ocm_copy", C.c_int, ocm_alloc_t, ocm_alloc_t, P(ocm_params))
_sig("ocm_copy_onesided", C.c_int, ocm_alloc_t, P(ocm_params))
_sig("ocm_strerror", C.c_char_p, C.c_int)
_sig("ocm_last_error", C.c_int)
_sig("ocm_rank", C.c_int)
_sig("ocm_num_ranks", C.c_int)
_sig("ocm_device", C.c_int)
_sig("ocm_remote_rank", C.c_int, ocm_alloc_t)
_sig("ocm_remote_ptr", C.c_void_p, ocm_alloc_t)
_sig("ocm_alloc_id", C.c_uint64, ocm_alloc_t)
_sig("ocm_slab_ptr", C.c_void_p, ocm_alloc_t, C.c_int)
_sig("ocm_slab_bytes", C.c_uint64, ocm_alloc_t)
_sig("ocm_multicast_ready", C.c_int)
_sig("ocm_prefetch_views", C.c_int, C.c_uint)
_sig("ocm_is_multicast", C.c_int, ocm_alloc_t)
_sig("ocm_copy_async", C.c_int, ocm_alloc_t, ocm_alloc_t, P(ocm_params), C.c_void_p)
_sig("ocm_copy_onesided_async", C.c_int, ocm_alloc_t, P(ocm_params), C.c_void_p)
_sig("ocm_quiet", C.c_int)
_sig("ocm_copy_onesided_batch", C.c_int, C.c_void_p, P(ocm_params), C.c_int, C.c_void_p)
_sig("ocm_signal_read", C.c_int, ocm_alloc_t, P(C.c_uint64))
_sig("ocm_signal_wait", C.c_int, ocm_alloc_t, C.c_uint64, C.c_int)
_sig("ocm_reduce_get_f32", C.c_int, ocm_alloc_t, P(ocm_params), C.c_void_p)
_sig("ocm_share", C.c_int, ocm_alloc_t, P(ocm_share_token))
_sig("ocm_attach", C.c_int, P(ocm_share_token), P(ocm_alloc_t))
_sig("ocm_attach_count", C.c_int, ocm_alloc_t)
_sig("ocm_copy_strided", C.c_int, ocm_alloc_t, ocm_alloc_t, P(ocm_strided), C.c_void_p)
_sig("ocm_copy_onesided_strided", C.c_int, ocm_alloc_t, P(ocm_strided), C.c_void_p)
_sig("ocm_get_stats", C.c_int, P(ocm_stats))
_sig("ocm_get_rank_stats", C.c_int, C.c_int, P(ocm_stats))
_sig("ocm_server_stats", C.c_int, P(C.c_uint64), P(C.c_uint64))
# kernel ABI
_sig("ocmk_copy", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p)
_sig("ocmk_copy_tuned", C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int,
     C.c_int, C.c_int, C.c_int, C.c_int)
_sig("ocmk_fill", C.c_int, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p)
_sig("ocmk_verify", C.c_int, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_void_p)
_sig("ocmk_checksum", C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p)
_sig("ocmk_signal", C.c_int, C.c_void_p, C.c_void_p)
_sig("ocmk_strided_copy", C.c_int, P(ocmk_strided), C.c_void_p, C.c_void_p)
_sig("ocmk_put_replicated", C.c_int, P(C.c_void_p), C.c_int, C.c_void_p, C.c_size_t, C.c_void_p)
_sig("ocmk_gups", C.c_int, P(C.c_void_p), C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p)
_sig("ocmk_gups_threads", C.c_uint64, C.c_uint64)
_sig("ocmk_sort_u64", C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p)
_sig("ocmk_sort_u64_noalloc", C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p)
_sig("ocmk_launch_count", C.c_uint64)
# oracles
_sig("ocm_oracle_copy", None, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)
_sig("ocm_oracle_fill", None, C.c_void_p, C.c_size_t, C.c_uint64)
_sig("ocm_oracle_verify", C.c_uint64, C.c_void_p, C.c_size_t, C.c_uint64)
_sig("ocm_oracle_checksum", C.c_uint64, C.c_void_p, C.c_size_t)
_sig("ocm_oracle_strided_copy", C.c_int, P(ocmk_strided))
_sig("ocm_oracle_gups", None, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64)
_sig("ocm_oracle_sort_u64", None, C.c_void_p, C.c_size_t)
_sig("ocmk_merge_u64", C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p)
_sig("ocmk_merge_runs_u64", C.c_int, C.c_void_p, C.c_void_p, P(C.c_uint64), C.c_int, C.c_void_p)
_sig("ocm_oracle_f32_to_e4m3", C.c_uint8, C.c_float)
_sig("ocm_oracle_e4m3_to_f32", C.c_float, C.c_uint8)


def strerror(err: int) -> str:
    return lib.ocm_strerror(int(err)).decode()


class OcmError(RuntimeError):
    def __init__(self, what: str, err: int | None = None):
        err = lib.ocm_last_error() if err is None else abs(err)
        super().__init__(f"{what}: {strerror(err)} (ocm_err {err})")
        self.err = err

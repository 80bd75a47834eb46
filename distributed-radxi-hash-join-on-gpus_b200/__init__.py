"""B200-native radix hash join: Python launcher over the C ABI in include/rhj.h.

The product is `librhj_b200.so` (hand-written sm_100a CUDA behind an extern "C"
boundary).  This package only loads it with ctypes and offers a thin mirror of the
reference's operator surface (`HashJoin`, `Configuration`) for tests, the benchmark
and multi-process (torchrun) launches.  There is no CPU fallback: without the
compiled library, or without a GPU, every compute call raises.

torch is used for plumbing only: device buffers, streams and `torch.distributed`
(NCCL all-gather of the per-rank histograms, barriers).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "librhj_b200.so")
INCLUDE = os.path.join(ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


class RhjError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rhj error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    """rhj_config == the knobs of core/Configuration.h:19-38."""

    _fields_ = [("network_fanout_bits", C.c_uint32), ("local_fanout_bits", C.c_uint32),
                ("payload_bits", C.c_uint32), ("enable_two_level", C.c_uint32),
                ("tuple_bytes", C.c_uint32), ("build_capacity", C.c_uint32),
                ("allocation_factor", C.c_double), ("chunk_tuples", C.c_uint32), ("flags", C.c_uint32)]


FLAG_HIST_WARP_AGGREGATE = 1
FLAG_STAGE_HOST_INPUTS = 2
FLAG_BALANCED_ASSIGNMENT = 4
FLAG_DEBUG_ENV = 8
FLAG_STAGED_EXCHANGE = 16
FLAG_DIRECT_EXCHANGE = 32


def make_config(network_fanout_bits=5, local_fanout_bits=5, payload_bits=27, enable_two_level=False,
                tuple_bytes=16, build_capacity=0, allocation_factor=1.1, chunk_tuples=0, flags=0) -> Config:
    """Defaults are the reference's shipped values (core/Configuration.h:28-37)."""
    return Config(network_fanout_bits, local_fanout_bits, payload_bits, int(bool(enable_two_level)), tuple_bytes,
                  build_capacity, allocation_factor, chunk_tuples, flags)


class Result(C.Structure):
    _fields_ = [("matches", C.c_uint64), ("sum_key", C.c_uint64), ("sum_rid_r", C.c_uint64),
                ("sum_rid_s", C.c_uint64), ("sum_mix", C.c_uint64)]

    def astuple(self):
        return (self.matches, self.sum_key, self.sum_rid_r, self.sum_rid_s, self.sum_mix)


class Timings(C.Structure):
    _fields_ = [("jtotal_us", C.c_double), ("jhist_us", C.c_double), ("jmpi_us", C.c_double),
                ("jproc_us", C.c_double), ("lp_us", C.c_double), ("bp_us", C.c_double), ("h2d_us", C.c_double),
                ("snetcompl_us", C.c_double), ("push_us", C.c_double)]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class WindowHandle(C.Structure):
    _fields_ = [("ipc", C.c_ubyte * 64), ("inner_offset", C.c_uint64), ("outer_offset", C.c_uint64),
                ("capacity_inner", C.c_uint64), ("capacity_outer", C.c_uint64), ("flags_offset", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/rhj.h declares
_vp, _u64, _u32, _int = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
_cfgp, _resp = C.POINTER(Config), C.POINTER(Result)
SYMBOLS = {
    "rhj_last_error": (C.c_char_p, []),
    "rhj_version": (C.c_char_p, []),
    "rhj_launch_count": (_u64, []),
    "rhj_init": (_int, [C.POINTER(_vp), _int, _int, _int, _cfgp]),
    "rhj_destroy": (_int, [_vp]),
    "rhj_pool_allocate": (_int, [_vp, _u64]),
    "rhj_pool_get_memory": (_int, [_vp, _u64, C.POINTER(_vp)]),
    "rhj_pool_reset": (_int, [_vp]),
    "rhj_pool_free_all": (_int, [_vp]),
    "rhj_workspace_bytes": (_u64, [_cfgp, _int, _u64, _u64, _u64, _u64]),
    "rhj_plan": (_int, [_vp, _u64, _u64, _u64, _u64]),
    "rhj_window_export": (_int, [_vp, C.POINTER(WindowHandle)]),
    "rhj_window_import": (_int, [_vp, _int, C.POINTER(WindowHandle)]),
    "rhj_window_set_peer": (_int, [_vp, _int, _vp, _vp]),
    "rhj_exchange_has_flags": (_int, [_vp]),
    "rhj_exchange_is_staged": (_int, [_vp]),
    "rhj_window_get": (_int, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "rhj_histogram": (_int, [_vp, _vp, _u64, _vp, _u64, _vp]),
    "rhj_local_histogram_dev": (_vp, [_vp]),
    "rhj_histogram_rel": (_int, [_vp, _int, _vp, _u64, _vp]),
    "rhj_compute_offsets_rel": (_int, [_vp, _int, _vp, _vp]),
    "rhj_compute_offsets": (_int, [_vp, _vp, _vp]),
    "rhj_get_assignment": (_int, [_vp, _vp, _vp]),
    "rhj_get_local_histogram": (_int, [_vp, _int, _vp, _vp]),
    "rhj_get_global_histogram": (_int, [_vp, _int, _vp, _vp]),
    "rhj_get_base_offsets": (_int, [_vp, _int, _vp, _vp]),
    "rhj_get_write_offsets": (_int, [_vp, _int, _vp, _vp]),
    "rhj_network_partition": (_int, [_vp, _vp]),
    "rhj_network_partition_rel": (_int, [_vp, _int, _vp]),
    "rhj_set_sm_limit": (_int, [_vp, _int]),
    "rhj_local_partition": (_int, [_vp, _vp]),
    "rhj_exchange_wait": (_int, [_vp, _int, _vp]),
    "rhj_exchange_sync": (_int, [_vp, _int]),
    "rhj_local_partition_rel": (_int, [_vp, _int, _vp]),
    "rhj_get_local_partition": (_int, [_vp, _int, _u32, _vp, _vp, _vp]),
    "rhj_get_partition_dev": (_int, [_vp, _int, _u32, _int, C.POINTER(_vp), C.POINTER(_u64)]),
    "rhj_build_probe": (_int, [_vp, _vp]),
    "rhj_build_probe_accumulate": (_int, [_vp, _vp]),
    "rhj_build_probe_pairs": (_int, [_vp, _vp, _u64, _vp]),
    "rhj_get_result": (_int, [_vp, _resp, _vp]),
    "rhj_get_timings": (_int, [_vp, C.POINTER(Timings)]),
    "rhj_join": (_int, [_vp, _vp, _u64, _vp, _u64, _vp, _resp]),
    "rhj_join_host": (_int, [_vp, _vp, _u64, _vp, _u64, _vp, _resp]),
    "rhj_stage_host_inputs": (_int, [_vp, _vp, _u64, _vp, _u64, _vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "rhj_fill_perm": (_int, [_vp, _u32, _u64, _u64, _u64, _u64, _u64, _u64, _u64, _vp]),
    "rhj_apply_perm": (_int, [_vp, _u64, _u32, _u64, _u64, _vp]),
    "rhj_local_histogram_noalloc": (_int, [_vp, _u32, _u64, _u32, _vp, _vp]),
    "rhj_host_alloc": (_int, [C.POINTER(_vp), _u64]),
    "rhj_host_free": (_int, [_vp]),
    "rhj_memcpy_d2h": (_int, [_vp, _vp, _u64]),
    "rhj_memcpy_h2d": (_int, [_vp, _vp, _u64]),
    "rhj_host_assignment": (None, [_u32, _u32, _vp]),
    "rhj_host_offsets": (None, [_vp, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp]),
    "rhj_host_assignment_balanced": (None, [_vp, _vp, _u32, _u32, _vp]),
    "rhj_host_offsets_assigned": (None, [_vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp]),
}


def sources():
    return [os.path.join(CSRC, f) for f in ("rhj_api.cu", "rhj_kernels.cuh", "rhj_device.cuh")] + \
           [os.path.join(INCLUDE, "rhj.h")]


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile librhj_b200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    stale = (not os.path.exists(LIB_PATH)
             or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(s) for s in sources()))
    if force or stale:
        cmd = ["nvcc"] + NVCC_FLAGS + ["-o", LIB_PATH, os.path.join(CSRC, "rhj_api.cu")]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.check_call(cmd)
    return LIB_PATH


_lib = None


def lib():
    """The loaded C ABI.  Raises (never falls back) when the library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RhjError(-2, f"{LIB_PATH} is not built; run __graft_entry__.build() (nvcc, sm_100a). "
                               "There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int):
    if rc != 0:
        raise RhjError(rc, lib().rhj_last_error().decode())


def _ptr(x) -> int:
    """Device/host address of a torch tensor, numpy array or raw int."""
    if x is None:
        return 0
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    raise TypeError(type(x))


def _stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


class Context:
    """One rank's rhj_ctx (one per GPU).  Mirrors what HashJoin's constructor receives
    (operators/HashJoin.cpp:31-39: numberOfNodes, nodeId) plus the Configuration knobs."""

    def __init__(self, cfg: Config, rank: int = 0, nranks: int = 1, device: int = 0):
        self.cfg, self.rank, self.nranks, self.device = cfg, rank, nranks, device
        self.P = 1 << cfg.network_fanout_bits
        self.L = 1 << cfg.local_fanout_bits
        h = _vp()
        check(lib().rhj_init(C.byref(h), rank, nranks, device, C.byref(cfg)))
        self.h = h

    def sm_count(self) -> int:
        import torch
        return torch.cuda.get_device_properties(self.device).multi_processor_count

    def close(self):
        if getattr(self, "h", None):
            lib().rhj_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- pool / plan / windows
    def pool_allocate(self, nbytes: int):
        check(lib().rhj_pool_allocate(self.h, nbytes))

    def pool_get_memory(self, nbytes: int) -> int:
        out = _vp()
        check(lib().rhj_pool_get_memory(self.h, nbytes, C.byref(out)))
        return out.value

    def plan(self, max_inner: int, max_outer: int, window_inner_cap: int = 0, window_outer_cap: int = 0):
        check(lib().rhj_plan(self.h, max_inner, max_outer, window_inner_cap, window_outer_cap))

    def window_export(self) -> WindowHandle:
        wh = WindowHandle()
        check(lib().rhj_window_export(self.h, C.byref(wh)))
        return wh

    def window_import(self, peer: int, wh: WindowHandle):
        check(lib().rhj_window_import(self.h, peer, C.byref(wh)))

    def is_staged(self) -> bool:
        return bool(lib().rhj_exchange_is_staged(self.h))

    def has_flags(self) -> bool:
        return bool(lib().rhj_exchange_has_flags(self.h))

    def window_get(self):
        a, b = _vp(), _vp()
        check(lib().rhj_window_get(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def window_set_peer(self, peer: int, inner_ptr: int, outer_ptr: int):
        check(lib().rhj_window_set_peer(self.h, peer, inner_ptr, outer_ptr))

    # ---- stages (tasks/Task.h:10-30)
    def histogram(self, inner, inner_n: int, outer, outer_n: int):
        check(lib().rhj_histogram(self.h, _ptr(inner), inner_n, _ptr(outer), outer_n, _stream()))

    def histogram_rel(self, which: int, rel, n: int):
        check(lib().rhj_histogram_rel(self.h, which, _ptr(rel), n, _stream()))

    def compute_offsets_rel(self, which: int, gathered=None):
        check(lib().rhj_compute_offsets_rel(self.h, which, _ptr(gathered), _stream()))

    def local_histogram_ptr(self) -> int:
        return lib().rhj_local_histogram_dev(self.h)

    def compute_offsets(self, gathered=None):
        check(lib().rhj_compute_offsets(self.h, _ptr(gathered), _stream()))

    def network_partition(self):
        check(lib().rhj_network_partition(self.h, _stream()))

    def network_partition_rel(self, which: int):
        check(lib().rhj_network_partition_rel(self.h, which, _stream()))

    def set_sm_limit(self, sms: int):
        check(lib().rhj_set_sm_limit(self.h, sms))

    def local_partition(self):
        check(lib().rhj_local_partition(self.h, _stream()))

    def exchange_wait(self, which: int):
        check(lib().rhj_exchange_wait(self.h, which, _stream()))

    def exchange_sync(self, which: int):
        check(lib().rhj_exchange_sync(self.h, which))

    def local_partition_rel(self, which: int):
        check(lib().rhj_local_partition_rel(self.h, which, _stream()))

    def build_probe(self):
        check(lib().rhj_build_probe(self.h, _stream()))

    def build_probe_accumulate(self):
        check(lib().rhj_build_probe_accumulate(self.h, _stream()))

    def build_probe_pairs(self, pairs, capacity: int | None = None):
        """Build/probe with materialised output: `pairs` is a device int64 tensor with room for
        2 * capacity values, filled with (rid_R, rid_S) rows in unspecified order."""
        cap = capacity if capacity is not None else pairs.numel() // 2
        if pairs.numel() < 2 * cap:
            raise ValueError("pairs tensor is smaller than 2 * capacity")
        check(lib().rhj_build_probe_pairs(self.h, pairs.data_ptr(), cap, _stream()))

    def result(self) -> Result:
        r = Result()
        check(lib().rhj_get_result(self.h, C.byref(r), _stream()))
        return r

    def timings(self) -> Timings:
        t = Timings()
        check(lib().rhj_get_timings(self.h, C.byref(t)))
        return t

    # ---- getters (tasks/HistogramComputation.cpp:78-130)
    def _get_u64(self, fn, which: int):
        import numpy as np
        out = np.zeros(self.P, np.uint64)
        check(fn(self.h, which, out.ctypes.data, _stream()))
        return out

    def get_assignment(self):
        import numpy as np
        out = np.zeros(self.P, np.uint32)
        check(lib().rhj_get_assignment(self.h, out.ctypes.data, _stream()))
        return out

    def get_local_histogram(self, which):
        return self._get_u64(lib().rhj_get_local_histogram, which)

    def get_global_histogram(self, which):
        return self._get_u64(lib().rhj_get_global_histogram, which)

    def get_base_offsets(self, which):
        return self._get_u64(lib().rhj_get_base_offsets, which)

    def get_write_offsets(self, which):
        return self._get_u64(lib().rhj_get_write_offsets, which)

    def get_local_partition(self, which: int, p: int):
        import numpy as np
        hist = np.zeros(self.L, np.uint64)
        prefix = np.zeros(self.L, np.uint64)
        check(lib().rhj_get_local_partition(self.h, which, p, hist.ctypes.data, prefix.ctypes.data, _stream()))
        return hist, prefix

    def get_partition(self, which: int, p: int, after_local_pass: bool):
        """Host copy (numpy u64) of owned partition p: the window slice, or what the local pass wrote."""
        import numpy as np
        import torch
        ptr, n = _vp(), _u64()
        torch.cuda.current_stream().synchronize()
        check(lib().rhj_get_partition_dev(self.h, which, p, int(after_local_pass), C.byref(ptr), C.byref(n)))
        out = np.zeros(n.value, np.uint64)
        if n.value:
            check(lib().rhj_memcpy_d2h(out.ctypes.data, ptr.value, n.value * 8))
        return out

    # ---- the operator
    def join(self, inner, inner_n: int, outer, outer_n: int) -> Result:
        r = Result()
        check(lib().rhj_join(self.h, _ptr(inner), inner_n, _ptr(outer), outer_n, _stream(), C.byref(r)))
        return r

    def join_async(self, inner, inner_n: int, outer, outer_n: int):
        """All stages enqueued, no host synchronisation (result via .result())."""
        check(lib().rhj_join(self.h, _ptr(inner), inner_n, _ptr(outer), outer_n, _stream(), None))

    def stage_host_inputs(self, inner_host, inner_n: int, outer_host, outer_n: int):
        """Host relations -> the context's device staging buffers (asynchronous on the current stream);
        returns the two device addresses to hand to histogram()."""
        a, b = _vp(), _vp()
        check(lib().rhj_stage_host_inputs(self.h, _ptr(inner_host), inner_n, _ptr(outer_host), outer_n, _stream(),
                                          C.byref(a), C.byref(b)))
        return a.value or 0, b.value or 0

    def join_host(self, inner_host, inner_n: int, outer_host, outer_n: int) -> Result:
        r = Result()
        check(lib().rhj_join_host(self.h, _ptr(inner_host), inner_n, _ptr(outer_host), outer_n, _stream(), C.byref(r)))
        return r


class PinnedBuffer:
    """Page-locked host buffer (cudaHostAlloc) exposed as a numpy array."""

    def __init__(self, nbytes: int):
        import numpy as np
        p = _vp()
        check(lib().rhj_host_alloc(C.byref(p), nbytes))
        self.ptr, self.nbytes = p.value, nbytes
        self.array = np.ctypeslib.as_array((C.c_ubyte * nbytes).from_address(self.ptr))

    def free(self):
        if self.ptr:
            self.array = None
            lib().rhj_host_free(self.ptr)
            self.ptr = 0


def fill_perm(tensor, tuple_bytes: int, n: int, first: int, domain: int, seed: int, modulo: int = 0,
              start_key: int = 0, start_rid: int = 0):
    """Device generator, bit-identical to oracle.orc_fill_perm (harness-owned; Relation.cpp:63-85 stand-in)."""
    check(lib().rhj_fill_perm(_ptr(tensor), tuple_bytes, n, first, domain, seed, modulo, start_key, start_rid, _stream()))


def apply_perm(tensor, n: int, stride_words: int, domain: int, seed: int):
    """tensor[i * stride_words] <- perm(tensor[i * stride_words]) with fill_perm's bijection of (domain, seed)."""
    check(lib().rhj_apply_perm(_ptr(tensor), n, stride_words, domain, seed, _stream()))


from .hashjoin import HashJoin, HotSplitJoin, Relation, distribute, host_pieces, join_multi_rank, join_probe_pieces  # noqa: E402,F401

"""ctypes front for oracle/liboracle.so (the plain-C CPU oracle).

TEST INFRASTRUCTURE ONLY -- PARITY UNPINNED (see oracle/gpuprobe_oracle.h for
the provenance statement).  Importable only from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs; the product package
(gpu-operator_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")

GEN_UNIT, GEN_EDGE, GEN_BITS = 0, 1, 2
STREAM_A, STREAM_B = 0, 1


def build(force: bool = False) -> str:
    """Compile the C oracle in-tree (gcc, a second or so)."""
    src = os.path.join(HERE, "gpuprobe_oracle.c")
    hdr = os.path.join(HERE, "gpuprobe_oracle.h")
    stale = (
        force
        or not os.path.exists(LIB_PATH)
        or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr))
    )
    if stale:
        subprocess.run(["make", "-C", HERE, "-B" if force else "-s"], check=True,
                       stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        u32, u64, i32 = C.c_uint32, C.c_uint64, C.c_int
        fp = C.POINTER(C.c_float)
        L.oracle_mix32.restype = u32
        L.oracle_mix32.argtypes = [u32]
        L.oracle_hash.restype = u32
        L.oracle_hash.argtypes = [u32, u32, u64]
        L.oracle_gen_bits.restype = u32
        L.oracle_gen_bits.argtypes = [u32, u32, u64, i32]
        L.oracle_fill_inputs.restype = None
        L.oracle_fill_inputs.argtypes = [fp, fp, u64, u64, u32, i32]
        L.oracle_fill_inputs_mt.restype = None
        L.oracle_fill_inputs_mt.argtypes = [fp, fp, u64, u64, u32, i32, i32]
        L.oracle_vectoradd.restype = None
        L.oracle_vectoradd.argtypes = [fp, fp, fp, u64]
        L.oracle_vectoradd_mt.restype = None
        L.oracle_vectoradd_mt.argtypes = [fp, fp, fp, u64, i32]
        L.oracle_digest.restype = None
        L.oracle_digest.argtypes = [fp, u64, C.POINTER(u64), C.POINTER(u32)]
        L.oracle_vectoradd_digest.restype = None
        L.oracle_vectoradd_digest.argtypes = [u64, u64, u32, i32, i32, C.POINTER(u64), C.POINTER(u32)]
        L.oracle_sample_check.restype = u64
        L.oracle_sample_check.argtypes = [fp, fp, fp, u64]
        L.oracle_count_mismatch.restype = u64
        L.oracle_count_mismatch.argtypes = [fp, fp, u64]
        L.oracle_p2p_fill.restype = None
        L.oracle_p2p_fill.argtypes = [C.POINTER(u32), u64, u64, u32, i32]
        L.oracle_p2p_checksum.restype = None
        L.oracle_p2p_checksum.argtypes = [u64, u32, i32, i32, C.POINTER(u64), C.POINTER(u32)]
        L.oracle_allreduce_fill.restype = None
        L.oracle_allreduce_fill.argtypes = [C.POINTER(C.c_int32), u64, u64, u32, i32]
        L.oracle_allreduce_expected.restype = None
        L.oracle_allreduce_expected.argtypes = [C.POINTER(C.c_int32), u64, u64, u32, i32]
        L.oracle_allreduce_digest.restype = None
        L.oracle_allreduce_digest.argtypes = [u64, u64, u32, i32, i32, i32, C.POINTER(u64), C.POINTER(u32)]
        L.oracle_fp_env_ok.restype = i32
        L.oracle_fp_env_ok.argtypes = []
        _lib = L
    return _lib


def _fp(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_float))


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:  # pragma: no cover
        return max(1, os.cpu_count() or 1)


def hash32(seed: int, stream: int, i: int) -> int:
    return int(lib().oracle_hash(seed, stream, i))


def fill_inputs(n: int, seed: int = 0, kind: int = GEN_UNIT, first: int = 0, threads: int = 1):
    a = np.empty(n, dtype=np.float32)
    b = np.empty(n, dtype=np.float32)
    if threads > 1:
        lib().oracle_fill_inputs_mt(_fp(a), _fp(b), first, n, seed, kind, threads)
    else:
        lib().oracle_fill_inputs(_fp(a), _fp(b), first, n, seed, kind)
    return a, b


def vectoradd(a: np.ndarray, b: np.ndarray, threads: int = 1) -> np.ndarray:
    """The host loop c[i] = a[i] + b[i] (fp32, one rounding)."""
    assert a.shape == b.shape
    c = np.empty_like(a)
    if threads <= 1:
        lib().oracle_vectoradd(_fp(a), _fp(b), _fp(c), a.size)
    else:
        lib().oracle_vectoradd_mt(_fp(a), _fp(b), _fp(c), a.size, threads)
    return c


def vectoradd_into(a: np.ndarray, b: np.ndarray, c: np.ndarray, threads: int = 1) -> None:
    lib().oracle_vectoradd_mt(_fp(a), _fp(b), _fp(c), a.size, threads)


def digest(c: np.ndarray):
    s, x = C.c_uint64(0), C.c_uint32(0)
    lib().oracle_digest(_fp(c), c.size, C.byref(s), C.byref(x))
    return int(s.value), int(x.value)


def vectoradd_digest(n: int, seed: int = 0, kind: int = GEN_UNIT, first: int = 0, threads: int = 0):
    s, x = C.c_uint64(0), C.c_uint32(0)
    lib().oracle_vectoradd_digest(first, n, seed, kind, threads or host_threads(), C.byref(s), C.byref(x))
    return int(s.value), int(x.value)


def sample_check(a, b, c) -> int:
    return int(lib().oracle_sample_check(_fp(a), _fp(b), _fp(c), a.size))


def count_mismatch(x: np.ndarray, y: np.ndarray) -> int:
    assert x.size == y.size
    return int(lib().oracle_count_mismatch(_fp(x), _fp(y), x.size))


def p2p_fill(nwords: int, seed: int, src: int, first: int = 0) -> np.ndarray:
    w = np.empty(nwords, dtype=np.uint32)
    lib().oracle_p2p_fill(w.ctypes.data_as(C.POINTER(C.c_uint32)), first, nwords, seed, src)
    return w


def p2p_checksum(nwords: int, seed: int, src: int, threads: int = 0):
    s, x = C.c_uint64(0), C.c_uint32(0)
    lib().oracle_p2p_checksum(nwords, seed, src, threads or host_threads(), C.byref(s), C.byref(x))
    return int(s.value), int(x.value)


def allreduce_fill(n: int, seed: int, rank: int, first: int = 0) -> np.ndarray:
    x = np.empty(n, dtype=np.int32)
    lib().oracle_allreduce_fill(x.ctypes.data_as(C.POINTER(C.c_int32)), first, n, seed, rank)
    return x


def allreduce_expected(n: int, seed: int, nranks: int, first: int = 0) -> np.ndarray:
    x = np.empty(n, dtype=np.int32)
    lib().oracle_allreduce_expected(x.ctypes.data_as(C.POINTER(C.c_int32)), first, n, seed, nranks)
    return x


def allreduce_digest(n: int, seed: int, nranks: int, rounds: int = 1, first: int = 0, threads: int = 0):
    """(sum64, xor32) of what every rank holds after `rounds` in-place int32 sum-allreduces."""
    s, x = C.c_uint64(0), C.c_uint32(0)
    lib().oracle_allreduce_digest(first, n, seed, nranks, rounds, threads or host_threads(), C.byref(s), C.byref(x))
    return int(s.value), int(x.value)


def fp_env_ok() -> bool:
    return bool(lib().oracle_fp_env_ok())

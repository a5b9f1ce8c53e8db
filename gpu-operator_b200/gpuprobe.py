"""ctypes mirror of include/gpuprobe.h -- the Python-side stand-in for the cgo binding.

The reference's host code is Go and no Go toolchain exists on either box, so the calls a
Go `validator/gpuprobe.go` would make through cgo are mirrored here 1:1:
same entry points, same structs, same status codes.  Tests and bench.py call the library
only through this module, i.e. through the C ABI.

There is no fallback: if libgpuprobe.so has not been built, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# GPUPROBE_LIB selects another build of the same library (e.g. lib/libgpuprobe_ncclstatic.so, the variant the validator
# image ships); the default is the in-tree build.
LIB_PATH = os.environ.get("GPUPROBE_LIB") or os.path.join(HERE, "lib", "libgpuprobe.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "gpuprobe.h")

ABI_VERSION = 2

OK = 0
FAIL_VERIFY = 1
FAIL_BANDWIDTH = 2
EINVAL, ENODEV, EARCH, ECUDA, ENOMEM, ENOPEER, ENCCL = -1, -2, -3, -4, -5, -6, -7

GEN_UNIT, GEN_EDGE, GEN_BITS = 0, 1, 2

VADD_AUTO = 0
VADD_LDG128_U2, VADD_LDG128_U4, VADD_LDG128_U8 = 1, 2, 3
VADD_LDG256_U2, VADD_LDG256_U4 = 4, 5
VADD_TMA, VADD_PERSIST128, VADD_SCALAR, VADD_TMA_8K, VADD_LDG256_U2_EF = 6, 7, 8, 9, 10
VADD_LDG128_U1, VADD_LDG256_U1 = 11, 12
VADD_SMALL = 13
VADD_NUM_VARIANTS = 14
VARIANT_NAMES = {
    0: "auto", 1: "ldg128_u2", 2: "ldg128_u4", 3: "ldg128_u8", 4: "ldg256_u2", 5: "ldg256_u4",
    6: "tma_16k", 7: "persist128", 8: "scalar", 9: "tma_8k", 10: "ldg256_u2_ef", 11: "ldg128_u1", 12: "ldg256_u1", 13: "small",
}

P2P_SERIAL, P2P_ROUNDS, P2P_ROUNDS_PUSH, P2P_SERIAL_PUSH, P2P_SERIAL_CE, P2P_ROUNDS_CE = 0, 1, 2, 3, 4, 5


class DevInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("cc_major", C.c_int), ("cc_minor", C.c_int), ("sm_count", C.c_int),
                ("mem_clock_khz", C.c_int), ("mem_bus_bits", C.c_int), ("max_smem_optin", C.c_int),
                ("hbm_bytes", C.c_uint64), ("l2_bytes", C.c_uint64)]


class VaddCfg(C.Structure):
    _fields_ = [("dev", C.c_int), ("n_elems", C.c_uint64), ("iters", C.c_int), ("warmup", C.c_int),
                ("seed", C.c_uint32), ("gen_kind", C.c_int), ("variant", C.c_int), ("verify", C.c_int),
                ("peak_nominal_gbps", C.c_double), ("peak_measured_gbps", C.c_double)]


class VaddResult(C.Structure):
    _fields_ = [("n", C.c_uint64), ("sum64", C.c_uint64), ("xor32", C.c_uint32), ("variant", C.c_uint32),
                ("mismatches", C.c_uint64), ("algo_bytes", C.c_uint64),
                ("ms_best", C.c_double), ("ms_median", C.c_double), ("ms_mean", C.c_double),
                ("gbps_best", C.c_double), ("gbps_median", C.c_double),
                ("frac_nominal", C.c_double), ("frac_measured", C.c_double),
                ("l2_resident", C.c_int), ("launches", C.c_int)]


class P2PCfg(C.Structure):
    _fields_ = [("ndev", C.c_int), ("devs", C.POINTER(C.c_int)), ("bytes_per_pair", C.c_uint64),
                ("iters", C.c_int), ("warmup", C.c_int), ("mode", C.c_int), ("seed", C.c_uint32),
                ("peak_gbps", C.c_double)]


class P2PCell(C.Structure):
    _fields_ = [("src", C.c_int), ("dst", C.c_int), ("peer_ok", C.c_int), ("verified", C.c_int),
                ("checksum_expected", C.c_uint64), ("checksum_got", C.c_uint64),
                ("xor_expected", C.c_uint32), ("xor_got", C.c_uint32),
                ("ms_best", C.c_double), ("ms_median", C.c_double), ("gbps", C.c_double),
                ("frac_of_peak", C.c_double)]


class ArCfg(C.Structure):
    _fields_ = [("ndev", C.c_int), ("devs", C.POINTER(C.c_int)), ("count", C.c_uint64), ("iters", C.c_int),
                ("warmup", C.c_int), ("seed", C.c_uint32), ("peak_gbps", C.c_double)]


class ArResult(C.Structure):
    _fields_ = [("ndev", C.c_int), ("nccl_version", C.c_int), ("bytes", C.c_uint64), ("mismatches", C.c_uint64),
                ("ms_best", C.c_double), ("ms_median", C.c_double), ("algbw_gbps", C.c_double),
                ("busbw_gbps", C.c_double), ("frac_of_peak", C.c_double),
                ("sum64", C.c_uint64), ("xor32", C.c_uint32), ("rounds", C.c_uint32), ("ranks_equal", C.c_int),
                ("impl", C.c_int)]


AR_IMPL_NCCL, AR_IMPL_NVLS, AR_IMPL_P2P, AR_IMPL_HYBRID, AR_IMPL_PUSH, AR_IMPL_CE, AR_IMPL_NVLS_CE = 0, 1, 2, 3, 4, 5, 6
AR_IMPL_NAMES = {0: "nccl", 1: "nvls", 2: "two_shot", 3: "hybrid", 4: "push", 5: "copy_engine", 6: "nvls_ce"}


class GpuProbeError(RuntimeError):
    """A gpuprobe entry point returned non-zero (code > 0: verification failed; < 0: could not run)."""

    def __init__(self, code: int, where: str, message: str):
        self.code = code
        self.where = where
        self.message = message
        super().__init__(f"{where}: [{code}] {strerror(code)}" + (f": {message}" if message else ""))


_u64, _u32, _int = C.c_uint64, C.c_uint32, C.c_int
_fp = C.POINTER(C.c_float)
_err = (C.c_char_p, C.c_size_t)

# name -> (restype, argtypes): every symbol include/gpuprobe.h declares
SIGNATURES = {
    "gpuprobe_abi_version": (_int, []),
    "gpuprobe_device_count": (_int, [C.POINTER(_int)]),
    "gpuprobe_device_info": (_int, [_int, C.POINTER(DevInfo)]),
    "gpuprobe_strerror": (C.c_char_p, [_int]),
    "gpuprobe_vectoradd": (_int, [C.POINTER(VaddCfg), C.POINTER(VaddResult), *_err]),
    "gpuprobe_vectoradd_buffers": (_int, [_int, C.c_void_p, C.c_void_p, C.c_void_p, _u64, *_err]),
    "gpuprobe_vectoradd_device": (_int, [_int, C.c_void_p, C.c_void_p, C.c_void_p, _u64, _int, C.c_void_p, *_err]),
    "gpuprobe_fill_inputs": (None, [C.c_void_p, C.c_void_p, _u64, _u64, _u32, _int]),
    "gpuprobe_fill_inputs_device": (_int, [_int, C.c_void_p, C.c_void_p, _u64, _u64, _u32, _int, C.c_void_p, *_err]),
    "gpuprobe_digest_device": (_int, [_int, C.c_void_p, _u64, C.POINTER(_u64), C.POINTER(_u32), C.c_void_p, *_err]),
    "gpuprobe_digest_device_noalloc": (_int, [_int, C.c_void_p, _u64, C.POINTER(_u64), C.POINTER(_u32), C.c_void_p, *_err]),
    "gpuprobe_verify_device": (_int, [_int, C.c_void_p, _u64, _u64, _u32, _int, C.POINTER(_u64), C.c_void_p, *_err]),
    "gpuprobe_host_alloc": (_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "gpuprobe_host_free": (_int, [C.c_void_p]),
    "gpuprobe_p2p_matrix": (_int, [C.POINTER(P2PCfg), C.POINTER(P2PCell), *_err]),
    "gpuprobe_p2p_loopback": (_int, [_int, _u64, _int, _int, _u32, C.POINTER(P2PCell), *_err]),
    "gpuprobe_p2p_expected": (None, [_u64, _u32, _int, _int, C.POINTER(_u64), C.POINTER(_u32)]),
    "gpuprobe_allreduce_i32": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_nvls": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_p2p": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_hybrid": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_push": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_auto": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_ce": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_allreduce_i32_nvls_ce": (_int, [C.POINTER(ArCfg), C.POINTER(ArResult), *_err]),
    "gpuprobe_shutdown": (None, []),
}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()' or "
        f"make -C gpu-operator_b200/csrc). There is no CPU fallback.")

lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = the library does not export what the header declares
    _fn.restype = _res
    _fn.argtypes = _args

if lib.gpuprobe_abi_version() != ABI_VERSION:
    raise ImportError(f"libgpuprobe ABI {lib.gpuprobe_abi_version()} != binding ABI {ABI_VERSION}")


def strerror(code: int) -> str:
    return lib.gpuprobe_strerror(code).decode()


def _check(code: int, where: str, errbuf) -> None:
    if code != OK:
        raise GpuProbeError(code, where, errbuf.value.decode(errors="replace"))


def _errbuf():
    return C.create_string_buffer(512)


def _ptr(x) -> int:
    """Address of a numpy array / torch tensor / ctypes pointer / int."""
    if x is None:
        return 0
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return int(x.data_ptr())
    if hasattr(x, "ctypes"):
        return int(x.ctypes.data)
    return C.cast(x, C.c_void_p).value or 0


def device_count() -> int:
    n = _int(0)
    rc = lib.gpuprobe_device_count(C.byref(n))
    if rc != OK:
        raise GpuProbeError(rc, "gpuprobe_device_count", "")
    return n.value


def device_info(dev: int = 0) -> DevInfo:
    info = DevInfo()
    rc = lib.gpuprobe_device_info(dev, C.byref(info))
    if rc != OK:
        raise GpuProbeError(rc, "gpuprobe_device_info", "")
    return info


def vectoradd(n: int, dev: int = 0, iters: int = 20, warmup: int = 3, seed: int = 0, gen_kind: int = GEN_UNIT,
              variant: int = VADD_AUTO, verify: int = 3, peak_nominal_gbps: float = 0.0,
              peak_measured_gbps: float = 0.0) -> VaddResult:
    cfg = VaddCfg(dev, n, iters, warmup, seed, gen_kind, variant, verify, peak_nominal_gbps, peak_measured_gbps)
    out, eb = VaddResult(), _errbuf()
    _check(lib.gpuprobe_vectoradd(C.byref(cfg), C.byref(out), eb, len(eb)), "gpuprobe_vectoradd", eb)
    return out


def vectoradd_buffers(a, b, c, n: int, dev: int = 0) -> None:
    eb = _errbuf()
    _check(lib.gpuprobe_vectoradd_buffers(dev, _ptr(a), _ptr(b), _ptr(c), n, eb, len(eb)),
           "gpuprobe_vectoradd_buffers", eb)


def vectoradd_device(d_a, d_b, d_c, n: int, dev: int = 0, variant: int = VADD_AUTO, stream: int = 0) -> None:
    eb = _errbuf()
    _check(lib.gpuprobe_vectoradd_device(dev, _ptr(d_a), _ptr(d_b), _ptr(d_c), n, variant, stream, eb, len(eb)),
           "gpuprobe_vectoradd_device", eb)


def fill_inputs(a, b, n: int, seed: int = 0, gen_kind: int = GEN_UNIT, first: int = 0) -> None:
    lib.gpuprobe_fill_inputs(_ptr(a), _ptr(b), first, n, seed, gen_kind)


def fill_inputs_device(d_a, d_b, n: int, dev: int = 0, seed: int = 0, gen_kind: int = GEN_UNIT, first: int = 0,
                       stream: int = 0) -> None:
    eb = _errbuf()
    _check(lib.gpuprobe_fill_inputs_device(dev, _ptr(d_a), _ptr(d_b), first, n, seed, gen_kind, stream, eb, len(eb)),
           "gpuprobe_fill_inputs_device", eb)


def digest_device(d_words, n: int, dev: int = 0, stream: int = 0):
    s, x, eb = _u64(0), _u32(0), _errbuf()
    _check(lib.gpuprobe_digest_device(dev, _ptr(d_words), n, C.byref(s), C.byref(x), stream, eb, len(eb)),
           "gpuprobe_digest_device", eb)
    return s.value, x.value


NOALLOC_MAX_WORDS = 1 << 30


def digest_device_noalloc(d_words, n: int, dev: int = 0, stream: int = 0):
    """gpuprobe_digest_device_noalloc: same result as digest_device, but no device allocation
    (fixed in-library scratch; at most NOALLOC_MAX_WORDS words per call)."""
    s, x, eb = _u64(0), _u32(0), _errbuf()
    _check(lib.gpuprobe_digest_device_noalloc(dev, _ptr(d_words), n, C.byref(s), C.byref(x), stream, eb, len(eb)),
           "gpuprobe_digest_device_noalloc", eb)
    return s.value, x.value


def verify_device(d_c, n: int, dev: int = 0, seed: int = 0, gen_kind: int = GEN_UNIT, first: int = 0,
                  stream: int = 0) -> int:
    bad, eb = _u64(0), _errbuf()
    _check(lib.gpuprobe_verify_device(dev, _ptr(d_c), first, n, seed, gen_kind, C.byref(bad), stream, eb, len(eb)),
           "gpuprobe_verify_device", eb)
    return bad.value


def host_alloc(nbytes: int) -> int:
    p = C.c_void_p(0)
    rc = lib.gpuprobe_host_alloc(C.byref(p), nbytes)
    if rc != OK:
        raise GpuProbeError(rc, "gpuprobe_host_alloc", f"{nbytes} bytes")
    return p.value


def host_free(p: int) -> None:
    lib.gpuprobe_host_free(p)


def pinned_array(n: int, dtype="float32"):
    """numpy view over pinned host memory from gpuprobe_host_alloc; keep the returned owner alive."""
    import numpy as np

    dt = np.dtype(dtype)
    addr = host_alloc(max(1, n) * dt.itemsize)
    buf = (C.c_char * (n * dt.itemsize)).from_address(addr)
    arr = np.frombuffer(buf, dtype=dt, count=n)

    class _Owner:
        def __init__(self, a):
            self.addr = a

        def free(self):
            if self.addr:
                host_free(self.addr)
                self.addr = 0

    return arr, _Owner(addr)


def p2p_expected(nwords: int, seed: int, src: int, threads: int = 0):
    s, x = _u64(0), _u32(0)
    lib.gpuprobe_p2p_expected(nwords, seed, src, threads, C.byref(s), C.byref(x))
    return s.value, x.value


def p2p_loopback(nbytes: int, dev: int = 0, iters: int = 3, kernel: int = 0, seed: int = 0) -> P2PCell:
    """gpuprobe_p2p_loopback: the P2P copy+checksum kernels with both buffers on one device (kernel 0 = bulk ring, 1 = ldst)."""
    cell, eb = P2PCell(), _errbuf()
    _check(lib.gpuprobe_p2p_loopback(dev, nbytes, iters, kernel, seed, C.byref(cell), eb, len(eb)), "gpuprobe_p2p_loopback", eb)
    return cell


def p2p_matrix(ndev: int, bytes_per_pair: int, iters: int = 10, warmup: int = 3, mode: int = P2P_SERIAL,
               seed: int = 0, devs=None, peak_gbps: float = 0.0, raise_on_fail: bool = True):
    """Returns (rc, cells) with cells[src_index][dst_index]."""
    arr = (C.c_int * ndev)(*devs) if devs is not None else None
    cfg = P2PCfg(ndev, arr, bytes_per_pair, iters, warmup, mode, seed, peak_gbps)
    cells = (P2PCell * (ndev * ndev))()
    eb = _errbuf()
    rc = lib.gpuprobe_p2p_matrix(C.byref(cfg), cells, eb, len(eb))
    if raise_on_fail:
        _check(rc, "gpuprobe_p2p_matrix", eb)
    return rc, [[cells[s * ndev + t] for t in range(ndev)] for s in range(ndev)]


def _allreduce(symbol: str, ndev: int, count: int, iters: int, warmup: int, seed: int, devs, peak_gbps: float) -> ArResult:
    arr = (C.c_int * ndev)(*devs) if devs is not None else None
    cfg = ArCfg(ndev, arr, count, iters, warmup, seed, peak_gbps)
    out, eb = ArResult(), _errbuf()
    _check(getattr(lib, symbol)(C.byref(cfg), C.byref(out), eb, len(eb)), symbol, eb)
    return out


def allreduce_i32(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                  peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32: ncclAllReduce(int32, sum) bracketed by the library's fill / verify / digest kernels."""
    return _allreduce("gpuprobe_allreduce_i32", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_nvls(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                       peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_nvls: the hand-written multimem (NVLS) allreduce kernel."""
    return _allreduce("gpuprobe_allreduce_i32_nvls", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_p2p(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                      peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_p2p: two-shot allreduce over peer mappings (2, 4 or 8 ranks)."""
    return _allreduce("gpuprobe_allreduce_i32_p2p", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_hybrid(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                         peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_hybrid: NVLS for the head of each slice, peer mappings for the tail."""
    return _allreduce("gpuprobe_allreduce_i32_hybrid", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_push(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                       peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_push: two-shot with posted writes only (scatter to the owner, add, broadcast)."""
    return _allreduce("gpuprobe_allreduce_i32_push", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_ce(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                     peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_ce: two-shot with the copy engines moving the data and a kernel adding."""
    return _allreduce("gpuprobe_allreduce_i32_ce", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_nvls_ce(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                          peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_nvls_ce: the multimem kernel on the head of the buffer, the copy-engine two-shot on the tail."""
    return _allreduce("gpuprobe_allreduce_i32_nvls_ce", ndev, count, iters, warmup, seed, devs, peak_gbps)


def allreduce_i32_auto(ndev: int, count: int, iters: int = 20, warmup: int = 5, seed: int = 0, devs=None,
                       peak_gbps: float = 0.0) -> ArResult:
    """gpuprobe_allreduce_i32_auto: the library's own choice of hand-written kernel for this rank count."""
    return _allreduce("gpuprobe_allreduce_i32_auto", ndev, count, iters, warmup, seed, devs, peak_gbps)


ALLREDUCE_IMPLS = {"nccl": allreduce_i32, "nvls": allreduce_i32_nvls, "two_shot": allreduce_i32_p2p,
                   "hybrid": allreduce_i32_hybrid, "push": allreduce_i32_push, "copy_engine": allreduce_i32_ce,
                   "nvls_ce": allreduce_i32_nvls_ce, "auto": allreduce_i32_auto}


def shutdown() -> None:
    lib.gpuprobe_shutdown()

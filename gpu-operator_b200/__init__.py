"""gpu-operator_b200 -- the GPU Operator validator's CUDA-workload path, rebuilt for B200.

Holds only what that one path needs:
  csrc/        hand-written sm_100a CUDA kernels + the C ABI (include/gpuprobe.h) -> lib/libgpuprobe.so
  gpuprobe.py  ctypes mirror of the C ABI (what a cgo binding would be; Go is absent on both boxes)

The directory name has a hyphen (it mirrors the reference repo's name), so it is loaded with
importlib under the module name `gpu_operator_b200` -- see __graft_entry__.load_package().
"""
from . import gpuprobe  # noqa: F401  (raises ImportError if libgpuprobe.so is not built: no fallback)

__all__ = ["gpuprobe"]

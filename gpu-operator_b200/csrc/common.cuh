// common.cuh -- shared host-side plumbing of libgpuprobe: error reporting, per-device
// state (serialising mutex, streams, events), small PTX wrappers.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/gpuprobe.h"

namespace gp {

constexpr int MAX_DEVICES = 64;

inline void set_err(char *err, size_t errlen, const char *fmt, ...) {
    if (!err || errlen == 0) return;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, errlen, fmt, ap);
    va_end(ap);
}

// Evaluates a CUDA runtime call; on failure writes the message and returns the mapped code
// from the enclosing function. cudaErrorMemoryAllocation maps to GPUPROBE_ENOMEM.
#define GP_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            gp::set_err(err, errlen, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            (void)cudaGetLastError();                                                              \
            return e_ == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;             \
        }                                                                                          \
    } while (0)

// One staging slot of the host-buffer path (gpuprobe_vectoradd_buffers): device chunks of A, B, C, pinned bounce
// buffers for pageable callers, and the event that says the slot's C has landed. Cached per device, grown on demand.
constexpr int BUF_MAX_SLOTS = 4;  // = number of copy streams per device
struct BufSlot {
    float *d_a = nullptr, *d_b = nullptr, *d_c = nullptr;
    float *h_a = nullptr, *h_b = nullptr, *h_c = nullptr;
    cudaEvent_t done = nullptr;
    uint64_t dev_cap = 0, host_cap = 0;       // elements the device / pinned buffers can hold
    uint64_t pending_off = 0, pending_n = 0;  // chunk whose C still sits in h_c
};

struct DeviceState {
    std::mutex mu;  // the synchronous probes (vectoradd, vectoradd_buffers, p2p, allreduce) on one device are serialised
    std::atomic<bool> inited{false};
    int dev = -1;
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
    size_t l2_bytes = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;       // compute / timing stream of the probes
    cudaStream_t copy_streams[BUF_MAX_SLOTS] = {};   // pipelined H2D / kernel / D2H chunks of the buffers path
    std::once_flag func_attrs_once;
    cudaError_t func_attrs_err = cudaSuccess;
    // Cached device scratch of the digest and verify kernels (partials + result + a mismatch counter). Allocated at the
    // first digest/verify on the device, kept until gpuprobe_shutdown; scratch_mu serialises its users, whichever stream
    // they run on (each of them synchronises its stream before it releases the lock).
    std::mutex scratch_mu;
    void *scratch = nullptr;
    // Cached staging of the host-buffer path; guarded by `mu`.
    BufSlot buf_slots[BUF_MAX_SLOTS];
};

DeviceState &device_state(int dev);

// Selects `dev`, checks it is an sm_100 part and lazily creates the cached state.
// Returns GPUPROBE_OK or an error code with the message in err.
int use_device(int dev, DeviceState **out, char *err, size_t errlen);

// Frees the cached staging of the host-buffer path (device current, st.mu held). Defined in api.cu.
void free_buf_slots(DeviceState &st);

// Shared launchers (defined in vectoradd.cu).
int launch_vadd(DeviceState &st, const float *a, const float *b, float *c, uint64_t n, int variant,
                cudaStream_t stream, char *err, size_t errlen);
int launch_fill_words(DeviceState &st, uint32_t *d_w, uint64_t n, uint32_t hash_base_value, cudaStream_t stream,
                      char *err, size_t errlen);
int run_digest(DeviceState &st, const void *d_words, uint64_t n, uint64_t *sum64, uint32_t *xor32,
               cudaStream_t stream, char *err, size_t errlen);

inline double median_of(std::vector<float> v) {
    if (v.empty()) return 0.0;
    std::sort(v.begin(), v.end());
    const size_t m = v.size() / 2;
    return (v.size() & 1u) ? v[m] : 0.5 * (static_cast<double>(v[m - 1]) + v[m]);
}

// ---- device-side helpers -----------------------------------------------------------
#if defined(__CUDACC__)

__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint32_t warp_xor_u32(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 128-bit streaming load through the read-only path, not allocated in L1.
__device__ __forceinline__ uint4 ldg128_stream(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg128_stream(void *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

#endif  // __CUDACC__

}  // namespace gp

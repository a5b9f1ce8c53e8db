// noalloc.cu -- gpuprobe_digest_device_noalloc: the digest of a device buffer for callers
// that must not allocate.
//
// Who that is.  A node validator does not always run on an idle GPU.  A health re-check that
// fires while tenant jobs are running meets two problems with the ordinary
// gpuprobe_digest_device(): (1) the jobs may hold practically all of HBM (or the whole MIG
// slice), so even a 14 KiB cudaMalloc can fail or eat into their headroom; (2) cudaMalloc and
// cudaFree take device-wide locks and cudaFree waits for the whole device to go idle, which
// stalls every tenant stream for the duration -- a probe must not do that to the workload it
// is checking on.  The same goes for callers that live inside somebody else's memory pool (a
// framework that owns all device memory and hands out slices of it) and are forbidden from
// going behind its back to the driver allocator.
//
// What it does instead.  The scratch the digest kernel needs (one partial sum/xor per CTA, a
// ticket counter, the result) lives in a fixed array in device global memory that is part of
// the library image: it exists on every GPU that loads libgpuprobe, from the moment the
// module is loaded until the process exits, and costs GPUPROBE_NOALLOC_SCRATCH_BYTES (192.25 KiB)
// per GPU whether or not this entry point is ever called.  Inputs and outputs are still the
// caller's.
//
// Consequences of a shared, permanent array:
//   * Size limit.  Every CTA digests a fixed 65,536-word segment and owns one partial slot, and
//     there are 16,384 slots, so at most GPUPROBE_NOALLOC_MAX_WORDS = 2^30 words (4 GiB) per
//     call.  Larger inputs are rejected with GPUPROBE_EINVAL; digest them in pieces (sums add
//     mod 2^64, xors xor) or use gpuprobe_digest_device.
//   * The array keeps whatever the previous call left in it.  The only state that must start
//     at zero is the ticket counter (and the result words); each call clears those 16 bytes on
//     the caller's stream before the kernel.  Partial slots need no clearing: every slot a
//     call reads was written by that same call.
//   * Two calls of one process that overlap on the same GPU (different streams, different
//     host threads) would race on the array and return garbage.  The library therefore
//     serialises calls per device on the host, and each call waits for its stream before it
//     returns, so calls made through this entry point never overlap each other.  Separate
//     processes do not interfere: each has its own copy of the array.
//   * The first call on a device resolves the array's device address (set-up, once per device).
//     Make that first call outside any stream capture -- and early: under CUDA's default lazy
//     module loading the address look-up can be the moment the library's device image (this
//     array included) is placed on the GPU, so a caller that waits until memory is exhausted
//     may see the set-up itself fail.  A call with n = 0 at start-up does the set-up and nothing else.
#include "common.cuh"

namespace gp {

constexpr int NA_THREADS = 256;
constexpr uint64_t NA_WORDS_PER_CTA = 65536;                                    // 256 KiB of input per CTA
constexpr uint64_t NA_MAX_PARTIALS = GPUPROBE_NOALLOC_MAX_WORDS / NA_WORDS_PER_CTA;  // 16,384
constexpr size_t NA_HEAD_BYTES = 16;  // uint64 sum, uint32 xor, uint32 ticket
static_assert(NA_HEAD_BYTES + NA_MAX_PARTIALS * (sizeof(uint64_t) + sizeof(uint32_t)) <= GPUPROBE_NOALLOC_SCRATCH_BYTES,
              "scratch array too small for the stated maximum");

// The scratch itself.  In HOST code this name is only a handle for cudaGetSymbolAddress -- it
// is not a device address and must never be passed to a kernel.
static __device__ __align__(256) unsigned char g_noalloc_scratch[GPUPROBE_NOALLOC_SCRATCH_BYTES];

struct NaHead {
    uint64_t out_sum;
    uint32_t out_xor;
    unsigned int ticket;
};
static_assert(sizeof(NaHead) == NA_HEAD_BYTES, "head layout");

// Variant of digest_kernel (vectoradd.cu): same reduction, but one fixed-size segment per CTA
// instead of a grid-stride loop, so the number of partial slots is a function of n alone.
__global__ void __launch_bounds__(NA_THREADS)
digest_noalloc_kernel(const uint32_t *__restrict__ w, uint64_t n, int vec_ok, unsigned char *scratch) {
    NaHead *head = reinterpret_cast<NaHead *>(scratch);
    uint64_t *partial_sum = reinterpret_cast<uint64_t *>(scratch + NA_HEAD_BYTES);
    uint32_t *partial_xor = reinterpret_cast<uint32_t *>(partial_sum + NA_MAX_PARTIALS);

    const uint64_t lo = static_cast<uint64_t>(blockIdx.x) * NA_WORDS_PER_CTA;
    const uint64_t hi = min(lo + NA_WORDS_PER_CTA, n);
    uint64_t s = 0;
    uint32_t x = 0;
    uint64_t done = lo;
    if (vec_ok) {
        const uint64_t n4 = (hi - lo) / 4;  // lo is a multiple of 4, so w + lo stays 16-byte aligned
        for (uint64_t i = threadIdx.x; i < n4; i += NA_THREADS) {
            const uint4 v = ldg128_stream(w + lo + i * 4);
            s += static_cast<uint64_t>(v.x) + v.y + static_cast<uint64_t>(v.z) + v.w;
            x ^= v.x ^ v.y ^ v.z ^ v.w;
        }
        done = lo + n4 * 4;
    }
    for (uint64_t i = done + threadIdx.x; i < hi; i += NA_THREADS) {
        const uint32_t v = __ldg(w + i);
        s += v;
        x ^= v;
    }

    __shared__ uint64_t sh_s[NA_THREADS / 32];
    __shared__ uint32_t sh_x[NA_THREADS / 32];
    __shared__ bool is_last;
    s = warp_sum_u64(s);
    x = warp_xor_u32(x);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sh_s[warp] = s;
        sh_x[warp] = x;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t bs = 0;
        uint32_t bx = 0;
        for (int i = 0; i < NA_THREADS / 32; ++i) {
            bs += sh_s[i];
            bx ^= sh_x[i];
        }
        partial_sum[blockIdx.x] = bs;
        partial_xor[blockIdx.x] = bx;
        __threadfence();
        is_last = (atomicAdd(&head->ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {  // the last CTA to arrive folds this call's partials
        __threadfence();
        uint64_t ts = 0;
        uint32_t tx = 0;
        for (unsigned int i = threadIdx.x; i < gridDim.x; i += NA_THREADS) {
            ts += *(volatile uint64_t *)(partial_sum + i);
            tx ^= *(volatile uint32_t *)(partial_xor + i);
        }
        ts = warp_sum_u64(ts);
        tx = warp_xor_u32(tx);
        __syncthreads();
        if (lane == 0) {
            sh_s[warp] = ts;
            sh_x[warp] = tx;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint64_t bs = 0;
            uint32_t bx = 0;
            for (int i = 0; i < NA_THREADS / 32; ++i) {
                bs += sh_s[i];
                bx ^= sh_x[i];
            }
            head->out_sum = bs;
            head->out_xor = bx;
            head->ticket = 0;
        }
    }
}

// Per-device set-up state: the device address of g_noalloc_scratch on that GPU.
struct NaDevice {
    std::mutex mu;  // serialises calls on the device: they share the array
    unsigned char *scratch = nullptr;
};
static NaDevice g_na[MAX_DEVICES];

}  // namespace gp

using namespace gp;

extern "C" int gpuprobe_digest_device_noalloc(int dev, const void *d_words, uint64_t n, uint64_t *sum64,
                                              uint32_t *xor32, void *stream, char *err, size_t errlen) {
    if (!sum64 || !xor32 || (n && !d_words)) {
        set_err(err, errlen, "null pointer");
        return GPUPROBE_EINVAL;
    }
    if (n > GPUPROBE_NOALLOC_MAX_WORDS) {
        set_err(err, errlen,
                "n = %llu words exceeds GPUPROBE_NOALLOC_MAX_WORDS = %llu (the fixed scratch array holds one partial "
                "per 65536-word segment); split the buffer or use gpuprobe_digest_device",
                (unsigned long long)n, (unsigned long long)GPUPROBE_NOALLOC_MAX_WORDS);
        return GPUPROBE_EINVAL;
    }
    if (reinterpret_cast<uintptr_t>(d_words) & 3u) {
        set_err(err, errlen, "d_words must be 4-byte aligned");
        return GPUPROBE_EINVAL;
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        set_err(err, errlen, "no CUDA device available");
        return GPUPROBE_ENODEV;
    }
    if (dev < 0 || dev >= count || dev >= MAX_DEVICES) {
        set_err(err, errlen, "device %d out of range (0..%d)", dev, count - 1);
        return GPUPROBE_ENODEV;
    }
    GP_CUDA(cudaSetDevice(dev));
    NaDevice &nd = g_na[dev];
    std::lock_guard<std::mutex> lock(nd.mu);

    if (!nd.scratch) {
        // set-up, once per device, with that device current
        int major = 0;
        GP_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10) {
            set_err(err, errlen, "device %d is sm_%dx; libgpuprobe kernels are built for sm_100a only", dev, major);
            return GPUPROBE_EARCH;
        }
        unsigned char *p = nullptr;
        GP_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_scratch));
        if (!p) {
            set_err(err, errlen, "cudaGetSymbolAddress returned a null scratch address on device %d", dev);
            return GPUPROBE_ECUDA;
        }
        nd.scratch = p;
    }
    if (n == 0) {
        *sum64 = 0;
        *xor32 = 0;
        return GPUPROBE_OK;
    }

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // the array still holds the previous call's result and, if that call died mid-kernel, a stale ticket
    GP_CUDA(cudaMemsetAsync(nd.scratch, 0, NA_HEAD_BYTES, s));
    const unsigned grid = static_cast<unsigned>((n + NA_WORDS_PER_CTA - 1) / NA_WORDS_PER_CTA);
    const int vec_ok = (reinterpret_cast<uintptr_t>(d_words) & 15u) == 0;
    digest_noalloc_kernel<<<grid, NA_THREADS, 0, s>>>(static_cast<const uint32_t *>(d_words), n, vec_ok, nd.scratch);
    GP_CUDA(cudaGetLastError());
    NaHead host{};
    GP_CUDA(cudaMemcpyAsync(&host, nd.scratch, sizeof(host), cudaMemcpyDeviceToHost, s));
    GP_CUDA(cudaStreamSynchronize(s));
    *sum64 = host.out_sum;
    *xor32 = host.out_xor;
    return GPUPROBE_OK;
}

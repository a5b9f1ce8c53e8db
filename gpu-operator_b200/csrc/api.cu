// api.cu -- library-level entry points of libgpuprobe: version, device queries, error strings,
// the host-side generator, pinned host memory, cached per-device state.
#include <thread>

#include "common.cuh"
#include "gen.cuh"

namespace gp {

static DeviceState g_state[MAX_DEVICES];
static std::mutex g_init_mu;

DeviceState &device_state(int dev) { return g_state[dev]; }

static std::atomic<int> g_device_count{-1};  // cached after the first successful query (devices do not come and go)

int use_device(int dev, DeviceState **out, char *err, size_t errlen) {
    int count = g_device_count.load(std::memory_order_acquire);
    if (count < 0) {
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0) {
            set_err(err, errlen, "no CUDA device available: %s", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
            (void)cudaGetLastError();
            return GPUPROBE_ENODEV;
        }
        g_device_count.store(count, std::memory_order_release);
    }
    if (dev < 0 || dev >= count || dev >= MAX_DEVICES) {
        set_err(err, errlen, "device %d out of range (0..%d)", dev, count - 1);
        return GPUPROBE_ENODEV;
    }
    GP_CUDA(cudaSetDevice(dev));
    DeviceState &st = g_state[dev];
    if (st.inited.load(std::memory_order_acquire)) {  // the launch-bound path: no lock, no driver query
        *out = &st;
        return GPUPROBE_OK;
    }
    std::lock_guard<std::mutex> lock(g_init_mu);
    if (!st.inited.load(std::memory_order_relaxed)) {
        cudaDeviceProp p{};
        GP_CUDA(cudaGetDeviceProperties(&p, dev));
        if (p.major != 10) {
            // The fat binary holds sm_100a SASS only: refuse loudly instead of failing at launch.
            set_err(err, errlen, "device %d (%s) is sm_%d%d; libgpuprobe kernels are built for sm_100a only", dev, p.name,
                    p.major, p.minor);
            return GPUPROBE_EARCH;
        }
        st.dev = dev;
        st.sm_count = p.multiProcessorCount;
        st.cc_major = p.major;
        st.cc_minor = p.minor;
        st.l2_bytes = static_cast<size_t>(p.l2CacheSize);
        st.max_smem_optin = static_cast<int>(p.sharedMemPerBlockOptin);
        cudaStream_t made[1 + BUF_MAX_SLOTS] = {};
        for (int i = 0; i < 1 + BUF_MAX_SLOTS; ++i) {
            cudaError_t e = cudaStreamCreateWithFlags(&made[i], cudaStreamNonBlocking);
            if (e != cudaSuccess) {  // leave nothing half-built behind
                for (int j = 0; j < i; ++j) cudaStreamDestroy(made[j]);
                set_err(err, errlen, "cudaStreamCreateWithFlags failed: %s", cudaGetErrorString(e));
                (void)cudaGetLastError();
                return GPUPROBE_ECUDA;
            }
        }
        st.stream = made[0];
        for (int i = 0; i < BUF_MAX_SLOTS; ++i) st.copy_streams[i] = made[1 + i];
        st.inited.store(true, std::memory_order_release);
    }
    *out = &st;
    return GPUPROBE_OK;
}

void free_buf_slots(DeviceState &st) {
    for (auto &sl : st.buf_slots) {
        cudaFree(sl.d_a);
        cudaFree(sl.d_b);
        cudaFree(sl.d_c);
        if (sl.h_a) cudaFreeHost(sl.h_a);
        if (sl.h_b) cudaFreeHost(sl.h_b);
        if (sl.h_c) cudaFreeHost(sl.h_c);
        if (sl.done) cudaEventDestroy(sl.done);
        sl = BufSlot{};
    }
    (void)cudaGetLastError();
}

static int host_threads(int want) {
    if (want > 0) return want;
    unsigned hc = std::thread::hardware_concurrency();
    return hc ? static_cast<int>(hc) : 1;
}

}  // namespace gp

using namespace gp;

extern "C" int gpuprobe_abi_version(void) { return GPUPROBE_ABI_VERSION; }

extern "C" int gpuprobe_device_count(int *n) {
    if (!n) return GPUPROBE_EINVAL;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        *n = 0;
        return GPUPROBE_ENODEV;
    }
    *n = count;
    return GPUPROBE_OK;
}

extern "C" int gpuprobe_device_info(int dev, gpuprobe_devinfo *out) {
    if (!out) return GPUPROBE_EINVAL;
    std::memset(out, 0, sizeof(*out));
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || dev < 0 || dev >= count) {
        (void)cudaGetLastError();
        return GPUPROBE_ENODEV;
    }
    cudaDeviceProp p{};
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return GPUPROBE_ECUDA;
    }
    std::snprintf(out->name, sizeof(out->name), "%.127s", p.name);
    out->cc_major = p.major;
    out->cc_minor = p.minor;
    out->sm_count = p.multiProcessorCount;
    out->mem_clock_khz = p.memoryClockRate;
    out->mem_bus_bits = p.memoryBusWidth;
    out->max_smem_optin = static_cast<int>(p.sharedMemPerBlockOptin);
    out->hbm_bytes = p.totalGlobalMem;
    out->l2_bytes = static_cast<uint64_t>(p.l2CacheSize);
    return GPUPROBE_OK;
}

extern "C" const char *gpuprobe_strerror(int code) {
    switch (code) {
    case GPUPROBE_OK: return "ok";
    case GPUPROBE_FAIL_VERIFY: return "probe ran but its output failed verification";
    case GPUPROBE_FAIL_BANDWIDTH: return "probe verified but bandwidth is below the requested fraction of peak";
    case GPUPROBE_EINVAL: return "invalid argument";
    case GPUPROBE_ENODEV: return "no such CUDA device (or the driver could not be loaded)";
    case GPUPROBE_EARCH: return "device is not sm_100; kernels are built for sm_100a only";
    case GPUPROBE_ECUDA: return "CUDA runtime error";
    case GPUPROBE_ENOMEM: return "out of device or pinned host memory";
    case GPUPROBE_ENOPEER: return "peer access unavailable between two devices";
    case GPUPROBE_ENCCL: return "NCCL unavailable or failed";
    default: return "unknown gpuprobe status";
    }
}

extern "C" void gpuprobe_fill_inputs(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int gen_kind) {
    if (gen_kind < GPUPROBE_GEN_UNIT || gen_kind > GPUPROBE_GEN_BITS) gen_kind = GPUPROBE_GEN_UNIT;
    const uint32_t base_a = hash_base(seed, STREAM_A), base_b = hash_base(seed, STREAM_B);
    for (uint64_t j = 0; j < n; ++j) {
        if (a) {
            const uint32_t u = gen_bits(base_a, first + j, gen_kind, false);
            std::memcpy(&a[j], &u, 4);
        }
        if (b) {
            const uint32_t u = gen_bits(base_b, first + j, gen_kind, true);
            std::memcpy(&b[j], &u, 4);
        }
    }
}

extern "C" void gpuprobe_p2p_expected(uint64_t nwords, uint32_t seed, int src, int threads, uint64_t *sum64,
                                      uint32_t *xor32) {
    const int nt = static_cast<int>(std::min<uint64_t>(static_cast<uint64_t>(host_threads(threads)), nwords / 65536 + 1));
    std::vector<uint64_t> sums(nt, 0);
    std::vector<uint32_t> xors(nt, 0);
    const uint32_t base = hash_base(seed, stream_p2p(src));
    auto work = [&](int t) {
        const uint64_t lo = nwords * t / nt, hi = nwords * (t + 1) / nt;
        uint64_t s = 0;
        uint32_t x = 0;
        for (uint64_t i = lo; i < hi; ++i) {
            const uint32_t w = hash_at(base, i);
            s += w;
            x ^= w;
        }
        sums[t] = s;
        xors[t] = x;
    };
    if (nt == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t) th.emplace_back(work, t);
        for (auto &t : th) t.join();
    }
    uint64_t s = 0;
    uint32_t x = 0;
    for (int t = 0; t < nt; ++t) {
        s += sums[t];
        x ^= xors[t];
    }
    if (sum64) *sum64 = s;
    if (xor32) *xor32 = x;
}

extern "C" int gpuprobe_host_alloc(void **p, size_t bytes) {
    if (!p) return GPUPROBE_EINVAL;
    *p = nullptr;
    cudaError_t e = cudaHostAlloc(p, bytes, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return e == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;
    }
    return GPUPROBE_OK;
}

extern "C" int gpuprobe_host_free(void *p) {
    if (!p) return GPUPROBE_OK;
    if (cudaFreeHost(p) != cudaSuccess) {
        (void)cudaGetLastError();
        return GPUPROBE_ECUDA;
    }
    return GPUPROBE_OK;
}

extern "C" void gpuprobe_shutdown(void) {
    std::lock_guard<std::mutex> lock(g_init_mu);
    for (int d = 0; d < MAX_DEVICES; ++d) {
        DeviceState &st = g_state[d];
        if (!st.inited.load(std::memory_order_acquire)) continue;
        std::lock_guard<std::mutex> dl(st.mu);
        std::lock_guard<std::mutex> sl(st.scratch_mu);
        if (cudaSetDevice(d) == cudaSuccess) {
            cudaDeviceSynchronize();
            free_buf_slots(st);
            cudaFree(st.scratch);
            cudaStreamDestroy(st.stream);
            for (auto &s : st.copy_streams) cudaStreamDestroy(s);
        }
        (void)cudaGetLastError();
        st.scratch = nullptr;
        st.stream = nullptr;
        for (auto &s : st.copy_streams) s = nullptr;
        st.inited.store(false, std::memory_order_release);
    }
}

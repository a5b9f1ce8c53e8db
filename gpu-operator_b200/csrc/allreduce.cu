// allreduce.cu -- probe (c): NCCL-over-NVLink int32 allreduce fabric check.  New capability
// (the reference has no collective of any kind).
//
// One process, one communicator over all probed devices (ncclCommInitAll) -- the shape of the
// cuda workload pod, which sees every GPU of the node (reference
// validator/cuda-workload-validation.yaml:21-25).  Rank r contributes x_r[i] = (int32)hash(seed, AR(r), i);
// the allreduce is in place, so after k rounds every rank must hold n^(k-1) * sum_r x_r[i] in
// wraparound int32 arithmetic, which the verify kernel recomputes from the generator.
//
// NCCL is loaded at run time (GPUPROBE_NCCL_LIB, else libnccl.so.2): linking libnccl_static.a
// adds ~400 MB to the library; the validator image build can do that instead.
#include <dlfcn.h>
#include <nccl.h>

#include "ce_twoshot.cuh"
#include "common.cuh"
#include "gen.cuh"

namespace gp {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int load_nccl(char *err, size_t errlen) {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.handle) return GPUPROBE_OK;
#if defined(GPUPROBE_NCCL_STATIC)
    // `make NCCL_STATIC=1`: libnccl_static.a is linked in (the validator image ships no CUDA user libraries).
    (void)err;
    (void)errlen;
    g_nccl.GetVersion = &ncclGetVersion;
    g_nccl.CommInitAll = &ncclCommInitAll;
    g_nccl.CommDestroy = &ncclCommDestroy;
    g_nccl.AllReduce = &ncclAllReduce;
    g_nccl.GroupStart = &ncclGroupStart;
    g_nccl.GroupEnd = &ncclGroupEnd;
    g_nccl.GetErrorString = &ncclGetErrorString;
    g_nccl.handle = &g_nccl;
    return GPUPROBE_OK;
#endif
    // Order: an explicit path; a copy some other component of this process already loaded (two different NCCLs
    // in one process break whichever is bound second); the system library.
    void *h = nullptr;
    const char *explicit_path = getenv("GPUPROBE_NCCL_LIB");
    if (explicit_path && *explicit_path) h = dlopen(explicit_path, RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
    for (const char *c : {"libnccl.so.2", "libnccl.so"}) {
        if (h) break;
        h = dlopen(c, RTLD_NOW | RTLD_LOCAL);
    }
    if (!h) {
        set_err(err, errlen, "cannot load NCCL (set GPUPROBE_NCCL_LIB): %s", dlerror());
        return GPUPROBE_ENCCL;
    }
    NcclApi api;
    api.handle = h;
#define GP_SYM(field, name)                                                     \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, name));         \
    if (!api.field) {                                                           \
        set_err(err, errlen, "NCCL library lacks symbol %s", name);            \
        dlclose(h);                                                             \
        return GPUPROBE_ENCCL;                                                  \
    }
    GP_SYM(GetVersion, "ncclGetVersion")
    GP_SYM(CommInitAll, "ncclCommInitAll")
    GP_SYM(CommDestroy, "ncclCommDestroy")
    GP_SYM(AllReduce, "ncclAllReduce")
    GP_SYM(GroupStart, "ncclGroupStart")
    GP_SYM(GroupEnd, "ncclGroupEnd")
    GP_SYM(GetErrorString, "ncclGetErrorString")
#undef GP_SYM
    g_nccl = api;
    return GPUPROBE_OK;
}

// Counts elements of x that differ from mult * sum_r hash(seed, AR(r), i) (uint32 wraparound).
__global__ void __launch_bounds__(256)
ar_verify_kernel(const uint32_t *__restrict__ x, uint64_t n, uint32_t seed, int nranks, uint32_t mult,
                 unsigned long long *mismatches) {
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    unsigned long long bad = 0;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t s = 0;
        for (int r = 0; r < nranks; ++r) s += hash(seed, stream_ar(r), i);
        bad += (s * mult != x[i]);
    }
    bad = warp_sum_u64(bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatches, bad);
}

struct ArDev {
    int ord = -1;
    DeviceState *st = nullptr;
    uint32_t *buf = nullptr;
    unsigned long long *bad = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

}  // namespace gp

using namespace gp;

extern "C" int gpuprobe_allreduce_i32(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    if (!cfg || !out) {
        set_err(err, errlen, "null cfg/out");
        return GPUPROBE_EINVAL;
    }
    std::memset(out, 0, sizeof(*out));
    const int nd = cfg->ndev;
    if (nd < 2 || nd > MAX_DEVICES || cfg->count == 0 || cfg->iters < 1 || cfg->warmup < 0) {
        set_err(err, errlen, "bad allreduce cfg: ndev=%d count=%llu iters=%d warmup=%d", nd,
                (unsigned long long)cfg->count, cfg->iters, cfg->warmup);
        return GPUPROBE_EINVAL;
    }
    int rc = load_nccl(err, errlen);
    if (rc) return rc;

    std::vector<ArDev> D(nd);
    std::vector<int> ords(nd);
    for (int i = 0; i < nd; ++i) {
        D[i].ord = ords[i] = cfg->devs ? cfg->devs[i] : i;
        for (int j = 0; j < i; ++j)
            if (ords[j] == ords[i]) {
                set_err(err, errlen, "device %d listed twice", ords[i]);
                return GPUPROBE_EINVAL;
            }
        rc = use_device(D[i].ord, &D[i].st, err, errlen);
        if (rc) return rc;
    }
    std::vector<int> order(nd);
    for (int i = 0; i < nd; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return D[x].ord < D[y].ord; });
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int i : order) locks.emplace_back(D[i].st->mu);

    std::vector<ncclComm_t> comms(nd, nullptr);
    bool comms_live = false;
    auto cleanup = [&]() {
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaDeviceSynchronize();
            cudaFree(d.buf);
            cudaFree(d.bad);
            if (d.ev0) cudaEventDestroy(d.ev0);
            if (d.ev1) cudaEventDestroy(d.ev1);
        }
        if (comms_live)
            for (auto c : comms)
                if (c) g_nccl.CommDestroy(c);
        (void)cudaGetLastError();
    };
#define GP_TRY(call)                                                                                   \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            set_err(err, errlen, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            (void)cudaGetLastError();                                                                  \
            cleanup();                                                                                 \
            return e_ == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;                 \
        }                                                                                              \
    } while (0)
#define GP_NCCL(call)                                                                                  \
    do {                                                                                               \
        ncclResult_t r_ = (call);                                                                      \
        if (r_ != ncclSuccess) {                                                                       \
            set_err(err, errlen, "%s failed: %s", #call, g_nccl.GetErrorString(r_));                   \
            cleanup();                                                                                 \
            return GPUPROBE_ENCCL;                                                                     \
        }                                                                                              \
    } while (0)

    GP_NCCL(g_nccl.GetVersion(&out->nccl_version));
    const uint64_t bytes = cfg->count * sizeof(int32_t);
    out->ndev = nd;
    out->bytes = bytes;
    for (int r = 0; r < nd; ++r) {
        ArDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaMalloc(&d.buf, bytes));
        GP_TRY(cudaMalloc(&d.bad, sizeof(unsigned long long)));
        GP_TRY(cudaMemsetAsync(d.bad, 0, sizeof(unsigned long long), d.st->stream));
        GP_TRY(cudaEventCreate(&d.ev0));
        GP_TRY(cudaEventCreate(&d.ev1));
        rc = launch_fill_words(*d.st, d.buf, cfg->count, hash_base(cfg->seed, stream_ar(r)), d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
    }
    GP_NCCL(g_nccl.CommInitAll(comms.data(), nd, ords.data()));
    comms_live = true;

    std::vector<float> ms;
    const int total = cfg->warmup + cfg->iters;
    for (int it = 0; it < total; ++it) {
        const bool timed = it >= cfg->warmup;
        if (timed)
            for (auto &d : D) {
                GP_TRY(cudaSetDevice(d.ord));
                GP_TRY(cudaEventRecord(d.ev0, d.st->stream));
            }
        GP_NCCL(g_nccl.GroupStart());
        for (int r = 0; r < nd; ++r)
            GP_NCCL(g_nccl.AllReduce(D[r].buf, D[r].buf, cfg->count, ncclInt32, ncclSum, comms[r], D[r].st->stream));
        GP_NCCL(g_nccl.GroupEnd());
        float worst = 0;
        for (auto &d : D) {
            GP_TRY(cudaSetDevice(d.ord));
            if (timed) GP_TRY(cudaEventRecord(d.ev1, d.st->stream));
            GP_TRY(cudaStreamSynchronize(d.st->stream));
            if (timed) {
                float m = 0;
                GP_TRY(cudaEventElapsedTime(&m, d.ev0, d.ev1));
                worst = std::max(worst, m);
            }
        }
        if (timed) ms.push_back(worst);
    }

    // after `total` in-place rounds every element is n^(total-1) * sum_r x_r[i]  (mod 2^32)
    uint32_t mult = 1;
    for (int k = 1; k < total; ++k) mult *= static_cast<uint32_t>(nd);
    uint64_t bad_total = 0;
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        const uint64_t want = (cfg->count + 255) / 256;
        const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(d.st->sm_count) * 8)));
        ar_verify_kernel<<<grid, 256, 0, d.st->stream>>>(d.buf, cfg->count, cfg->seed, nd, mult, d.bad);
        GP_TRY(cudaGetLastError());
        unsigned long long h = 0;
        GP_TRY(cudaMemcpyAsync(&h, d.bad, sizeof(h), cudaMemcpyDeviceToHost, d.st->stream));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
        bad_total += h;
    }
    // The reduced buffers themselves, digested: what a caller compares with a value computed elsewhere (the tests use
    // the CPU oracle). `mismatches` above is derived from this library's own generator and is a sanity figure only.
    out->ranks_equal = 1;
    for (int r = 0; r < nd; ++r) {
        ArDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        uint64_t s64 = 0;
        uint32_t x32 = 0;
        rc = run_digest(*d.st, d.buf, cfg->count, &s64, &x32, d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
        if (r == 0) {
            out->sum64 = s64;
            out->xor32 = x32;
        } else if (s64 != out->sum64 || x32 != out->xor32) {
            out->ranks_equal = 0;
        }
    }
    out->rounds = static_cast<uint32_t>(total);
    out->impl = GPUPROBE_AR_IMPL_NCCL;
    out->mismatches = bad_total;
    out->ms_best = *std::min_element(ms.begin(), ms.end());
    out->ms_median = median_of(ms);
    out->algbw_gbps = out->ms_median > 0 ? bytes / (out->ms_median * 1e6) : 0;
    out->busbw_gbps = out->algbw_gbps * 2.0 * (nd - 1) / nd;
    out->frac_of_peak = out->busbw_gbps / (cfg->peak_gbps > 0 ? cfg->peak_gbps : 900.0);
    cleanup();
    if (bad_total) {
        set_err(err, errlen, "allreduce: %llu elements differ from the expected sum", (unsigned long long)bad_total);
        return GPUPROBE_FAIL_VERIFY;
    }
    if (!out->ranks_equal) {
        set_err(err, errlen, "allreduce: the ranks' buffers differ from one another");
        return GPUPROBE_FAIL_VERIFY;
    }
    return GPUPROBE_OK;
#undef GP_TRY
#undef GP_NCCL
}

// ---------------------------------------------------------------------------------------------------------------
// gpuprobe_allreduce_i32_ce: the copy-engine two-shot (ce_twoshot.cuh) over the whole buffer.
namespace gp {
struct CeDev {
    int ord = -1;
    DeviceState *st = nullptr;
    uint32_t *buf = nullptr, *stage = nullptr;
    unsigned long long *bad = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};
}  // namespace gp

extern "C" int gpuprobe_allreduce_i32_ce(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    if (!cfg || !out) {
        set_err(err, errlen, "null cfg/out");
        return GPUPROBE_EINVAL;
    }
    std::memset(out, 0, sizeof(*out));
    const int nd = cfg->ndev;
    if (nd < 2 || nd > MAX_DEVICES || cfg->count == 0 || cfg->iters < 1 || cfg->warmup < 0) {
        set_err(err, errlen, "bad allreduce cfg: ndev=%d count=%llu iters=%d warmup=%d", nd,
                (unsigned long long)cfg->count, cfg->iters, cfg->warmup);
        return GPUPROBE_EINVAL;
    }
    std::vector<CeDev> D(nd);
    int rc;
    for (int i = 0; i < nd; ++i) {
        D[i].ord = cfg->devs ? cfg->devs[i] : i;
        for (int j = 0; j < i; ++j)
            if (D[j].ord == D[i].ord) {
                set_err(err, errlen, "device %d listed twice", D[i].ord);
                return GPUPROBE_EINVAL;
            }
        rc = use_device(D[i].ord, &D[i].st, err, errlen);
        if (rc) return rc;
    }
    std::vector<int> order(nd);
    for (int i = 0; i < nd; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return D[x].ord < D[y].ord; });
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int i : order) locks.emplace_back(D[i].st->mu);

    const uint64_t count = cfg->count;
    CeTwoShot engine;
    auto cleanup = [&]() {
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaDeviceSynchronize();
        }
        engine.destroy();
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaFree(d.buf);
            cudaFree(d.stage);
            cudaFree(d.bad);
            if (d.ev0) cudaEventDestroy(d.ev0);
            if (d.ev1) cudaEventDestroy(d.ev1);
        }
        (void)cudaGetLastError();
    };
#define GP_TRY(call)                                                                                   \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            set_err(err, errlen, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            (void)cudaGetLastError();                                                                  \
            cleanup();                                                                                 \
            return e_ == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;                 \
        }                                                                                              \
    } while (0)

    // peer access both ways between every pair: the engines then copy directly over NVLink
    for (auto &x : D)
        for (auto &y : D) {
            if (x.ord == y.ord) continue;
            int can = 0;
            GP_TRY(cudaDeviceCanAccessPeer(&can, x.ord, y.ord));
            if (!can) {
                set_err(err, errlen, "device %d cannot access device %d", x.ord, y.ord);
                cleanup();
                return GPUPROBE_ENOPEER;
            }
            GP_TRY(cudaSetDevice(x.ord));
            cudaError_t e = cudaDeviceEnablePeerAccess(y.ord, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled)
                (void)cudaGetLastError();
            else
                GP_TRY(e);
        }
    const uint64_t bytes = count * sizeof(uint32_t);
    out->ndev = nd;
    out->bytes = bytes;
    std::vector<CeRank> ranks(nd);
    for (int r = 0; r < nd; ++r) {
        CeDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaMalloc(&d.buf, bytes));
        GP_TRY(cudaMalloc(&d.stage, CeTwoShot::stage_words(count, nd) * sizeof(uint32_t)));
        GP_TRY(cudaMalloc(&d.bad, sizeof(unsigned long long)));
        GP_TRY(cudaMemsetAsync(d.bad, 0, sizeof(unsigned long long), d.st->stream));
        GP_TRY(cudaEventCreate(&d.ev0));
        GP_TRY(cudaEventCreate(&d.ev1));
        rc = launch_fill_words(*d.st, d.buf, count, hash_base(cfg->seed, stream_ar(r)), d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
        ranks[r].ord = d.ord;
        ranks[r].sm_count = d.st->sm_count;
        ranks[r].buf = d.buf;
        ranks[r].stage = d.stage;
        ranks[r].main = d.st->stream;
    }
    GP_TRY(engine.init(ranks, count, 0));
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
    }

    std::vector<float> ms;
    const int total = std::max(1, cfg->warmup) + cfg->iters;
    for (int it = 0; it < total; ++it) {
        const bool timed = it >= total - cfg->iters;
        if (timed)
            for (auto &d : D) {
                GP_TRY(cudaSetDevice(d.ord));
                GP_TRY(cudaEventRecord(d.ev0, d.st->stream));
            }
        GP_TRY(engine.issue_round());
        float worst = 0;
        for (auto &d : D) {
            GP_TRY(cudaSetDevice(d.ord));
            if (timed) GP_TRY(cudaEventRecord(d.ev1, d.st->stream));
        }
        for (auto &d : D) {
            GP_TRY(cudaSetDevice(d.ord));
            GP_TRY(cudaStreamSynchronize(d.st->stream));
            if (timed) {
                float m = 0;
                GP_TRY(cudaEventElapsedTime(&m, d.ev0, d.ev1));
                worst = std::max(worst, m);
            }
        }
        if (timed) ms.push_back(worst);
    }

    uint32_t mult = 1;
    for (int k = 1; k < total; ++k) mult *= static_cast<uint32_t>(nd);
    uint64_t bad_total = 0;
    out->ranks_equal = 1;
    for (int r = 0; r < nd; ++r) {
        CeDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        const uint64_t want = (count + 255) / 256;
        const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(d.st->sm_count) * 8)));
        ar_verify_kernel<<<grid, 256, 0, d.st->stream>>>(d.buf, count, cfg->seed, nd, mult, d.bad);
        GP_TRY(cudaGetLastError());
        unsigned long long h = 0;
        GP_TRY(cudaMemcpyAsync(&h, d.bad, sizeof(h), cudaMemcpyDeviceToHost, d.st->stream));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
        bad_total += h;
        uint64_t s64 = 0;
        uint32_t x32 = 0;
        rc = run_digest(*d.st, d.buf, count, &s64, &x32, d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
        if (r == 0) {
            out->sum64 = s64;
            out->xor32 = x32;
        } else if (s64 != out->sum64 || x32 != out->xor32) {
            out->ranks_equal = 0;
        }
    }
    out->rounds = static_cast<uint32_t>(total);
    out->impl = GPUPROBE_AR_IMPL_CE;
    out->mismatches = bad_total;
    out->ms_best = *std::min_element(ms.begin(), ms.end());
    out->ms_median = median_of(ms);
    out->algbw_gbps = out->ms_median > 0 ? bytes / (out->ms_median * 1e6) : 0;
    out->busbw_gbps = out->algbw_gbps * 2.0 * (nd - 1) / nd;
    out->frac_of_peak = out->busbw_gbps / (cfg->peak_gbps > 0 ? cfg->peak_gbps : 900.0);
    cleanup();
    if (bad_total) {
        set_err(err, errlen, "allreduce (copy engines): %llu elements differ from the expected sum", (unsigned long long)bad_total);
        return GPUPROBE_FAIL_VERIFY;
    }
    if (!out->ranks_equal) {
        set_err(err, errlen, "allreduce (copy engines): the ranks' buffers differ from one another");
        return GPUPROBE_FAIL_VERIFY;
    }
    return GPUPROBE_OK;
#undef GP_TRY
}

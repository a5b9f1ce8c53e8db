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

// Copy-engine two-shot, the arithmetic step: x[i] += sum over the staged copies of the same words. All operands are in
// this GPU's HBM (the copy engines put them there); 128-bit accesses, grid-stride.
__global__ void __launch_bounds__(256)
ar_add_staged_kernel(uint32_t *__restrict__ x, const uint32_t *__restrict__ stage, uint64_t slot_stride, int nslots, uint64_t n) {
    const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    const uint64_t n4 = n / 4;  // x, stage and slot_stride are 16-byte aligned (chunks start on 32-word boundaries)
    for (uint64_t v = tid; v < n4; v += stride) {
        uint4 acc = *reinterpret_cast<const uint4 *>(x + v * 4);
        for (int s = 0; s < nslots; ++s) {
            const uint4 t = ldg128_stream(stage + static_cast<uint64_t>(s) * slot_stride + v * 4);
            acc.x += t.x;
            acc.y += t.y;
            acc.z += t.z;
            acc.w += t.w;
        }
        *reinterpret_cast<uint4 *>(x + v * 4) = acc;
    }
    for (uint64_t i = n4 * 4 + tid; i < n; i += stride) {
        uint32_t acc = x[i];
        for (int s = 0; s < nslots; ++s) acc += stage[static_cast<uint64_t>(s) * slot_stride + i];
        x[i] = acc;
    }
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
// Copy-engine two-shot (gpuprobe_allreduce_i32_ce). Measured on this fabric (profiles/r02_fabric.md): loads and stores
// issued by SMs cross NVLink as 128-byte packets (1.125 / 1.1875 link bytes per payload byte, 667-700 GB/s per direction
// with both directions busy), while the DMA engines move 770 GB/s per direction in the same situation. So this variant
// lets the copy engines do every transfer and keeps only the arithmetic in a kernel:
//   scatter  rank r copies its words of slice q into a staging slot in q's memory      (n-1 peer copies per chunk)
//   add      rank q adds the n-1 staged copies to its own words of slice q             (ar_add_staged_kernel, local HBM)
//   gather   rank q copies the reduced slice q into every rank's buffer                (n-1 peer copies per chunk)
// Slices are cut into chunks so that the add of chunk k runs while the engines move chunks k+1 (scatter) and k-1
// (gather); ordering is by CUDA events across the GPUs' streams (one process drives all GPUs), no kernel ever waits on
// another GPU. Peers are visited in rotated order so that every GPU receives from one sender at a time.
namespace gp {
struct CeDev {
    int ord = -1;
    DeviceState *st = nullptr;
    uint32_t *buf = nullptr, *stage = nullptr;
    unsigned long long *bad = nullptr;
    cudaStream_t s_scatter = nullptr, s_gather = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_go = nullptr, ev_gathered = nullptr;
    std::vector<cudaEvent_t> ev_scattered, ev_added;  // per chunk
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
    uint64_t per = (count + nd - 1) / nd;
    per = (per + 31) & ~31ull;  // slice q = [q*per, min((q+1)*per, count))
    auto slice_lo = [&](int q) { return std::min<uint64_t>(per * static_cast<uint64_t>(q), count); };
    auto slice_len = [&](int q) { return std::min<uint64_t>(slice_lo(q) + per, count) - slice_lo(q); };
    // chunks: enough of them that the add overlaps the transfers, few enough that a copy stays large
    int chunks = 2;  // measured at 2 ranks, 1 GiB: 1 chunk 635, 2 chunks 748, 4 731, 8 699, 16 647 GB/s (profiles/r02_fabric.md)
    if (const char *e = getenv("GPUPROBE_CE_CHUNKS")) chunks = std::max(1, std::min(64, atoi(e)));
    while (chunks > 1 && per / chunks < (1u << 18)) chunks /= 2;
    const uint64_t chunk_words = ((per + chunks - 1) / chunks + 31) & ~31ull;

    auto cleanup = [&]() {
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaDeviceSynchronize();
            cudaFree(d.buf);
            cudaFree(d.stage);
            cudaFree(d.bad);
            if (d.s_scatter) cudaStreamDestroy(d.s_scatter);
            if (d.s_gather) cudaStreamDestroy(d.s_gather);
            for (cudaEvent_t e : {d.ev0, d.ev1, d.ev_go, d.ev_gathered})
                if (e) cudaEventDestroy(e);
            for (cudaEvent_t e : d.ev_scattered)
                if (e) cudaEventDestroy(e);
            for (cudaEvent_t e : d.ev_added)
                if (e) cudaEventDestroy(e);
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
    for (int r = 0; r < nd; ++r) {
        CeDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaMalloc(&d.buf, bytes));
        GP_TRY(cudaMalloc(&d.stage, static_cast<size_t>(nd - 1) * per * sizeof(uint32_t)));
        GP_TRY(cudaMalloc(&d.bad, sizeof(unsigned long long)));
        GP_TRY(cudaMemsetAsync(d.bad, 0, sizeof(unsigned long long), d.st->stream));
        GP_TRY(cudaStreamCreateWithFlags(&d.s_scatter, cudaStreamNonBlocking));
        GP_TRY(cudaStreamCreateWithFlags(&d.s_gather, cudaStreamNonBlocking));
        GP_TRY(cudaEventCreate(&d.ev0));
        GP_TRY(cudaEventCreate(&d.ev1));
        GP_TRY(cudaEventCreateWithFlags(&d.ev_go, cudaEventDisableTiming));
        GP_TRY(cudaEventCreateWithFlags(&d.ev_gathered, cudaEventDisableTiming));
        d.ev_scattered.assign(chunks, nullptr);
        d.ev_added.assign(chunks, nullptr);
        for (auto &e : d.ev_scattered) GP_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto &e : d.ev_added) GP_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        rc = launch_fill_words(*d.st, d.buf, count, hash_base(cfg->seed, stream_ar(r)), d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
    }
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
    }

    // words [off, off+len) of chunk k inside a slice of `len_slice` words
    auto chunk_range = [&](uint64_t len_slice, int k, uint64_t *off, uint64_t *len) {
        *off = std::min<uint64_t>(chunk_words * static_cast<uint64_t>(k), len_slice);
        *len = std::min<uint64_t>(*off + chunk_words, len_slice) - *off;
    };
    // One in-place round, issued asynchronously on every GPU's streams.
    auto issue_round = [&](bool timed) -> cudaError_t {
        cudaError_t e;
        for (int r = 0; r < nd; ++r) {  // start: the main stream has waited for the whole previous round (see the end)
            CeDev &d = D[r];
            if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
            if (timed && (e = cudaEventRecord(d.ev0, d.st->stream)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(d.ev_go, d.st->stream)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(d.s_scatter, d.ev_go, 0)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(d.s_gather, d.ev_go, 0)) != cudaSuccess) return e;
        }
        for (int k = 0; k < chunks; ++k)  // scatter: chunk by chunk, every rank, peers in rotated order
            for (int r = 0; r < nd; ++r) {
                CeDev &d = D[r];
                if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
                for (int j = 1; j < nd; ++j) {
                    const int q = (r + j) % nd;
                    uint64_t off, len;
                    chunk_range(slice_len(q), k, &off, &len);
                    if (!len) continue;
                    const int slot = r < q ? r : r - 1;
                    e = cudaMemcpyPeerAsync(D[q].stage + static_cast<uint64_t>(slot) * per + off, D[q].ord, d.buf + slice_lo(q) + off, d.ord,
                                            len * sizeof(uint32_t), d.s_scatter);
                    if (e != cudaSuccess) return e;
                }
                if ((e = cudaEventRecord(d.ev_scattered[k], d.s_scatter)) != cudaSuccess) return e;
            }
        for (int k = 0; k < chunks; ++k)  // add, then gather the reduced chunk to everyone
            for (int q = 0; q < nd; ++q) {
                CeDev &d = D[q];
                if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
                for (int r = 0; r < nd; ++r)
                    if (r != q && (e = cudaStreamWaitEvent(d.st->stream, D[r].ev_scattered[k], 0)) != cudaSuccess) return e;
                uint64_t off, len;
                chunk_range(slice_len(q), k, &off, &len);
                if (len) {
                    const uint64_t want = (len / 4 + 255) / 256;
                    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(d.st->sm_count) * 8)));
                    ar_add_staged_kernel<<<grid, 256, 0, d.st->stream>>>(d.buf + slice_lo(q) + off, d.stage + off, per, nd - 1, len);
                    if ((e = cudaGetLastError()) != cudaSuccess) return e;
                }
                if ((e = cudaEventRecord(d.ev_added[k], d.st->stream)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(d.s_gather, d.ev_added[k], 0)) != cudaSuccess) return e;
                for (int j = 1; j < nd && len; ++j) {
                    const int p = (q + j) % nd;
                    e = cudaMemcpyPeerAsync(D[p].buf + slice_lo(q) + off, D[p].ord, d.buf + slice_lo(q) + off, d.ord, len * sizeof(uint32_t),
                                            d.s_gather);
                    if (e != cudaSuccess) return e;
                }
            }
        for (auto &d : D) {
            if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(d.ev_gathered, d.s_gather)) != cudaSuccess) return e;
        }
        for (auto &d : D) {  // end of the round on every main stream: every gather, incoming and outgoing, has landed
            if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
            for (auto &x : D)
                if ((e = cudaStreamWaitEvent(d.st->stream, x.ev_gathered, 0)) != cudaSuccess) return e;
            if (timed && (e = cudaEventRecord(d.ev1, d.st->stream)) != cudaSuccess) return e;
        }
        return cudaSuccess;
    };
    auto sync_all = [&]() -> cudaError_t {
        for (auto &d : D) {
            cudaError_t e = cudaSetDevice(d.ord);
            if (e == cudaSuccess) e = cudaStreamSynchronize(d.st->stream);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    };

    std::vector<float> ms;
    const int total = std::max(1, cfg->warmup) + cfg->iters;
    for (int it = 0; it < total; ++it) {
        const bool timed = it >= total - cfg->iters;
        GP_TRY(issue_round(timed));
        GP_TRY(sync_all());
        if (timed) {
            float worst = 0;
            for (auto &d : D) {
                float m = 0;
                GP_TRY(cudaEventElapsedTime(&m, d.ev0, d.ev1));
                worst = std::max(worst, m);
            }
            ms.push_back(worst);
        }
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

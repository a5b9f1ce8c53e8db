// p2p.cu -- probe (b): all-pairs NVLink copy + checksum.  New capability: the reference never
// touches a second GPU (its only kernel runs on device 0; SURVEY.md section 2 kernel inventory).
//
// For a directed pair src -> dst the kernel runs on dst and READS src's buffer through the peer
// mapping (the reads are the NVLink traffic: requests travel dst -> src, data returns src -> dst
// through NVSwitch), writes the words into a local buffer and reduces sum(uint32) -> uint64 and
// xor with warp shuffles.  The copy and the checksum are one pass: 1 x bytes over NVLink, 1 x
// bytes of local HBM writes, nothing re-read.  PUSH modes run the same kernel on src with the
// peer pointer on the store side.
//
// Default kernel (round 2): a cp.async.bulk ring.  One thread per CTA asks the SM's copy engine for
// 32 KiB at a time (global -> shared, mbarrier transaction count), all threads checksum the stage out
// of shared memory, the same thread sends it on with a bulk store.  Measured against the register
// path on a loaded fabric (both directions busy, profiles/r02_fabric.md): pull 667 vs 629 GB/s per
// direction, push 700 vs 695; one pair alone 776 vs 766.  The register-path kernel (every thread
// keeps 8 x 128-bit loads in flight) stays as env GPUPROBE_P2P_KERNEL=ldst.
//
// Remote latency is microseconds, so tens of MB must be in flight chip-wide: 2 CTAs per SM x 3
// stages x 32 KiB = 28 MB on 148 SMs.
#include <future>

#include "common.cuh"
#include "gen.cuh"

namespace gp {

constexpr int P2P_THREADS = 512;
constexpr int P2P_UNROLL = 8;
constexpr int P2P_MAX_BLOCKS = 148 * 4;

struct P2PScratch {
    uint64_t out_sum;
    uint32_t out_xor;
    unsigned int ticket;  // 0 at launch; the last CTA resets it
    uint64_t partial_sum[P2P_MAX_BLOCKS];
    uint32_t partial_xor[P2P_MAX_BLOCKS];
};

__global__ void __launch_bounds__(P2P_THREADS)
p2p_copy_checksum_kernel(const uint4 *__restrict__ src_peer, uint4 *__restrict__ dst_local, uint64_t n16,
                         P2PScratch *scr) {
    uint64_t s = 0;
    uint32_t x = 0;
    constexpr uint64_t TILE = static_cast<uint64_t>(P2P_THREADS) * P2P_UNROLL;
    const uint64_t n_tiles = n16 / TILE;
    for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const uint64_t base = t * TILE + threadIdx.x;
        uint4 v[P2P_UNROLL];
#pragma unroll
        for (int u = 0; u < P2P_UNROLL; ++u) v[u] = ldg128_stream(src_peer + base + u * P2P_THREADS);
#pragma unroll
        for (int u = 0; u < P2P_UNROLL; ++u) {
            stg128_stream(dst_local + base + u * P2P_THREADS, v[u]);
            s += static_cast<uint64_t>(v[u].x) + v[u].y + static_cast<uint64_t>(v[u].z) + v[u].w;
            x ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
        }
    }
    // ragged remainder (< one tile), grid-strided one vector at a time
    for (uint64_t i = n_tiles * TILE + static_cast<uint64_t>(blockIdx.x) * P2P_THREADS + threadIdx.x; i < n16;
         i += static_cast<uint64_t>(gridDim.x) * P2P_THREADS) {
        const uint4 v = ldg128_stream(src_peer + i);
        stg128_stream(dst_local + i, v);
        s += static_cast<uint64_t>(v.x) + v.y + static_cast<uint64_t>(v.z) + v.w;
        x ^= v.x ^ v.y ^ v.z ^ v.w;
    }

    __shared__ uint64_t sh_s[P2P_THREADS / 32];
    __shared__ uint32_t sh_x[P2P_THREADS / 32];
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
        for (int i = 0; i < P2P_THREADS / 32; ++i) {
            bs += sh_s[i];
            bx ^= sh_x[i];
        }
        scr->partial_sum[blockIdx.x] = bs;
        scr->partial_xor[blockIdx.x] = bx;
        __threadfence();
        is_last = (atomicAdd(&scr->ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        uint64_t ts = 0;
        uint32_t tx = 0;
        for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
            ts += *(volatile uint64_t *)(scr->partial_sum + i);
            tx ^= *(volatile uint32_t *)(scr->partial_xor + i);
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
            for (int i = 0; i < P2P_THREADS / 32; ++i) {
                bs += sh_s[i];
                bx ^= sh_x[i];
            }
            scr->out_sum = bs;
            scr->out_xor = bx;
            scr->ticket = 0;
        }
    }
}

// ---- the bulk-copy (TMA engine) ring ------------------------------------------------------------------
constexpr int P2PB_THREADS = 128;
constexpr int P2PB_CHUNK = 32768;  // bytes per bulk request
constexpr int P2PB_STAGES = 3;
constexpr int P2PB_CTAS_PER_SM = 2;
constexpr size_t P2PB_SMEM = static_cast<size_t>(P2PB_STAGES) * P2PB_CHUNK + P2PB_STAGES * sizeof(uint64_t);

__device__ __forceinline__ uint32_t p2p_smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// Iteration k of a CTA (stage s = k % STAGES): wait full[s]; every thread adds its share of the stage into its
// checksum registers (LDS.128); __syncthreads; thread 0 issues the bulk store of stage s and then refills the stage
// that iteration k-1 stored -- after cp.async.bulk.wait_group.read 1 has shown that store has finished reading it --
// with chunk k+STAGES-1. So STAGES-1 loads and up to 2 stores are in flight per CTA at any time.
__global__ void __launch_bounds__(P2PB_THREADS)
p2p_bulk_copy_checksum_kernel(const unsigned char *__restrict__ src, unsigned char *__restrict__ dst, uint64_t bytes,
                              P2PScratch *scr) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + static_cast<size_t>(P2PB_STAGES) * P2PB_CHUNK);
    const uint64_t n_chunks = bytes / P2PB_CHUNK;
    const uint64_t first = blockIdx.x, stride = gridDim.x;
    const uint64_t mine = first < n_chunks ? (n_chunks - first + stride - 1) / stride : 0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < P2PB_STAGES; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(p2p_smem_u32(&full[s])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto load = [&](uint64_t k) {  // thread 0 only
        const int s = static_cast<int>(k % P2PB_STAGES);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(p2p_smem_u32(&full[s])), "r"(P2PB_CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         p2p_smem_u32(smem + static_cast<size_t>(s) * P2PB_CHUNK)),
                     "l"(src + (first + k * stride) * P2PB_CHUNK), "r"(P2PB_CHUNK), "r"(p2p_smem_u32(&full[s]))
                     : "memory");
    };
    if (threadIdx.x == 0)
        for (uint64_t k = 0; k < static_cast<uint64_t>(P2PB_STAGES - 1) && k < mine; ++k) load(k);

    uint64_t sum = 0;
    uint32_t x = 0;
    for (uint64_t k = 0; k < mine; ++k) {
        const int s = static_cast<int>(k % P2PB_STAGES);
        const uint32_t parity = static_cast<uint32_t>((k / P2PB_STAGES) & 1u);
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "W_%=:\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
            "@!p bra W_%=;\n"
            "}\n" ::"r"(p2p_smem_u32(&full[s])),
            "r"(parity)
            : "memory");
        const uint4 *p = reinterpret_cast<const uint4 *>(smem + static_cast<size_t>(s) * P2PB_CHUNK);
#pragma unroll 4
        for (int i = threadIdx.x; i < P2PB_CHUNK / 16; i += P2PB_THREADS) {
            const uint4 v = p[i];
            sum += static_cast<uint64_t>(v.x) + v.y + static_cast<uint64_t>(v.z) + v.w;
            x ^= v.x ^ v.y ^ v.z ^ v.w;
        }
        __syncthreads();  // everyone has read stage s (and, one iteration ago, the stage about to be refilled)
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (first + k * stride) * P2PB_CHUNK),
                         "r"(p2p_smem_u32(smem + static_cast<size_t>(s) * P2PB_CHUNK)), "r"(P2PB_CHUNK)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (k + P2PB_STAGES - 1 < mine) {
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                load(k + P2PB_STAGES - 1);
            }
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");

    // ragged remainder (< one chunk), 16 bytes at a time through registers
    const uint64_t n16 = bytes / 16;
    const uint4 *src16 = reinterpret_cast<const uint4 *>(src);
    uint4 *dst16 = reinterpret_cast<uint4 *>(dst);
    for (uint64_t i = n_chunks * (P2PB_CHUNK / 16) + static_cast<uint64_t>(blockIdx.x) * P2PB_THREADS + threadIdx.x; i < n16;
         i += static_cast<uint64_t>(gridDim.x) * P2PB_THREADS) {
        const uint4 v = ldg128_stream(src16 + i);
        stg128_stream(dst16 + i, v);
        sum += static_cast<uint64_t>(v.x) + v.y + static_cast<uint64_t>(v.z) + v.w;
        x ^= v.x ^ v.y ^ v.z ^ v.w;
    }

    __shared__ uint64_t sh_s[P2PB_THREADS / 32];
    __shared__ uint32_t sh_x[P2PB_THREADS / 32];
    __shared__ bool is_last;
    sum = warp_sum_u64(sum);
    x = warp_xor_u32(x);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sh_s[warp] = sum;
        sh_x[warp] = x;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t bs = 0;
        uint32_t bx = 0;
        for (int i = 0; i < P2PB_THREADS / 32; ++i) {
            bs += sh_s[i];
            bx ^= sh_x[i];
        }
        scr->partial_sum[blockIdx.x] = bs;
        scr->partial_xor[blockIdx.x] = bx;
        __threadfence();
        is_last = (atomicAdd(&scr->ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {  // the last CTA to finish folds the partials
        __threadfence();
        uint64_t ts = 0;
        uint32_t tx = 0;
        for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
            ts += *(volatile uint64_t *)(scr->partial_sum + i);
            tx ^= *(volatile uint32_t *)(scr->partial_xor + i);
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
            for (int i = 0; i < P2PB_THREADS / 32; ++i) {
                bs += sh_s[i];
                bx ^= sh_x[i];
            }
            scr->out_sum = bs;
            scr->out_xor = bx;
            scr->ticket = 0;
        }
    }
}

struct P2PDev {
    int ord = -1;
    DeviceState *st = nullptr;
    uint32_t *src = nullptr;  // this device's source words (read by its peers)
    uint32_t *dst = nullptr;  // landing buffer for the copy this device performs
    P2PScratch *scr = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t exp_sum = 0;
    uint32_t exp_xor = 0;
};

static int p2p_bulk_grid(const DeviceState &st, uint64_t bytes) {
    const uint64_t chunks = bytes / P2PB_CHUNK + 1;
    const uint64_t cap = std::min<uint64_t>(static_cast<uint64_t>(st.sm_count) * P2PB_CTAS_PER_SM, P2P_MAX_BLOCKS);
    return static_cast<int>(std::max<uint64_t>(1, std::min(chunks, cap)));
}

static int p2p_grid(const DeviceState &st, uint64_t n16) {
    const uint64_t tiles = n16 / (static_cast<uint64_t>(P2P_THREADS) * P2P_UNROLL) + 1;
    const uint64_t cap = std::min<uint64_t>(static_cast<uint64_t>(st.sm_count) * 4, P2P_MAX_BLOCKS);
    return static_cast<int>(std::max<uint64_t>(1, std::min(tiles, cap)));
}

}  // namespace gp

using namespace gp;

extern "C" int gpuprobe_p2p_matrix(const gpuprobe_p2p_cfg *cfg, gpuprobe_p2p_cell *cells, char *err, size_t errlen) {
    if (!cfg || !cells) {
        set_err(err, errlen, "null cfg/cells");
        return GPUPROBE_EINVAL;
    }
    const int nd = cfg->ndev;
    if (nd < 2 || nd > MAX_DEVICES || cfg->iters < 1 || cfg->warmup < 0 || cfg->bytes_per_pair == 0 ||
        (cfg->bytes_per_pair & 15u) || cfg->mode < GPUPROBE_P2P_SERIAL || cfg->mode > GPUPROBE_P2P_ROUNDS_CE) {
        set_err(err, errlen, "bad p2p cfg: ndev=%d bytes=%llu iters=%d warmup=%d mode=%d", nd,
                (unsigned long long)cfg->bytes_per_pair, cfg->iters, cfg->warmup, cfg->mode);
        return GPUPROBE_EINVAL;
    }
    std::memset(cells, 0, sizeof(gpuprobe_p2p_cell) * nd * nd);
    const uint64_t bytes = cfg->bytes_per_pair, nwords = bytes / 4, n16 = bytes / 16;
    const double peak = cfg->peak_gbps > 0 ? cfg->peak_gbps : 900.0;

    std::vector<P2PDev> D(nd);
    for (int i = 0; i < nd; ++i) {
        D[i].ord = cfg->devs ? cfg->devs[i] : i;
        for (int j = 0; j < i; ++j)
            if (D[j].ord == D[i].ord) {
                set_err(err, errlen, "device %d listed twice", D[i].ord);
                return GPUPROBE_EINVAL;
            }
        int rc = use_device(D[i].ord, &D[i].st, err, errlen);
        if (rc) return rc;
    }
    // serialise against other probes on these devices; lock in ordinal order
    std::vector<int> order(nd);
    for (int i = 0; i < nd; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return D[x].ord < D[y].ord; });
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int i : order) locks.emplace_back(D[i].st->mu);

    // expected checksums on the host, overlapped with the GPU work
    std::vector<std::future<void>> expect;
    const int host_thr = std::max(1u, std::thread::hardware_concurrency() / static_cast<unsigned>(nd));
    for (int i = 0; i < nd; ++i)
        expect.emplace_back(std::async(std::launch::async, [&, i]() {
            gpuprobe_p2p_expected(nwords, cfg->seed, D[i].ord, host_thr, &D[i].exp_sum, &D[i].exp_xor);
        }));

    int rc = GPUPROBE_OK;
    auto cleanup = [&]() {
        for (auto &f : expect)
            if (f.valid()) f.wait();
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaDeviceSynchronize();
            cudaFree(d.src);
            cudaFree(d.dst);
            cudaFree(d.scr);
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

    // buffers, source words, peer mappings
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaMalloc(&d.src, bytes));
        GP_TRY(cudaMalloc(&d.dst, bytes));
        GP_TRY(cudaMalloc(&d.scr, sizeof(P2PScratch)));
        GP_TRY(cudaMemsetAsync(d.scr, 0, sizeof(P2PScratch), d.st->stream));
        GP_TRY(cudaEventCreate(&d.ev0));
        GP_TRY(cudaEventCreate(&d.ev1));
        rc = launch_fill_words(*d.st, d.src, nwords, hash_base(cfg->seed, stream_p2p(d.ord)), d.st->stream, err, errlen);
        if (rc) {
            cleanup();
            return rc;
        }
    }
    bool missing_peer = false;
    for (int s = 0; s < nd; ++s)
        for (int t = 0; t < nd; ++t) {
            gpuprobe_p2p_cell &c = cells[s * nd + t];
            c.src = D[s].ord;
            c.dst = D[t].ord;
            if (s == t) continue;
            int can = 0;
            GP_TRY(cudaDeviceCanAccessPeer(&can, D[t].ord, D[s].ord));
            c.peer_ok = can;
            if (!can) {
                missing_peer = true;
                continue;
            }
            GP_TRY(cudaSetDevice(D[t].ord));
            cudaError_t e = cudaDeviceEnablePeerAccess(D[s].ord, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                (void)cudaGetLastError();
            } else {
                GP_TRY(e);
            }
        }
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
    }

    // PULL (default): the kernel runs on dst and loads src's buffer over NVLink. PUSH: it runs on src, loads its own
    // buffer and stores into dst's landing buffer over NVLink (posted writes need no request packets on the way back).
    const char *kern_env = getenv("GPUPROBE_P2P_KERNEL");
    const bool bulk = !(kern_env && !strcmp(kern_env, "ldst"));
    if (bulk)
        for (auto &d : D) {
            GP_TRY(cudaSetDevice(d.ord));
            GP_TRY(cudaFuncSetAttribute(p2p_bulk_copy_checksum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        static_cast<int>(P2PB_SMEM)));
        }
    const bool push = cfg->mode == GPUPROBE_P2P_ROUNDS_PUSH || cfg->mode == GPUPROBE_P2P_SERIAL_PUSH;
    const bool serial = cfg->mode == GPUPROBE_P2P_SERIAL || cfg->mode == GPUPROBE_P2P_SERIAL_PUSH || cfg->mode == GPUPROBE_P2P_SERIAL_CE;
    // CE modes: the destination's copy engine moves the bytes; no kernel of ours is in the timed region
    const bool ce = cfg->mode == GPUPROBE_P2P_SERIAL_CE || cfg->mode == GPUPROBE_P2P_ROUNDS_CE;
    auto runner = [&](int s, int t) { return push ? s : t; };
    // One timed launch of src -> dst on the running device's stream (events recorded around it).
    auto launch_pair = [&](int s, int t, bool timed) -> cudaError_t {
        P2PDev &dd = D[runner(s, t)];
        cudaError_t e = cudaSetDevice(dd.ord);
        if (e != cudaSuccess) return e;
        if (timed && (e = cudaEventRecord(dd.ev0, dd.st->stream)) != cudaSuccess) return e;
        if (ce) {
            if ((e = cudaMemcpyPeerAsync(D[t].dst, D[t].ord, D[s].src, D[s].ord, bytes, dd.st->stream)) != cudaSuccess) return e;
        } else if (bulk)
            p2p_bulk_copy_checksum_kernel<<<p2p_bulk_grid(*dd.st, bytes), P2PB_THREADS, P2PB_SMEM, dd.st->stream>>>(
                reinterpret_cast<const unsigned char *>(D[s].src), reinterpret_cast<unsigned char *>(D[t].dst), bytes, dd.scr);
        else
            p2p_copy_checksum_kernel<<<p2p_grid(*dd.st, n16), P2P_THREADS, 0, dd.st->stream>>>(
                reinterpret_cast<const uint4 *>(D[s].src), reinterpret_cast<uint4 *>(D[t].dst), n16, dd.scr);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (timed) e = cudaEventRecord(dd.ev1, dd.st->stream);
        return e;
    };
    // After the pair's launches: read the kernel's checksum, digest the landed copy, fill the cell.
    auto finish_pair = [&](int s, int t, std::vector<float> &ms) -> int {
        P2PDev &rr = D[runner(s, t)];
        P2PDev &dd = D[t];
        gpuprobe_p2p_cell &c = cells[s * nd + t];
        P2PScratch head{};
        GP_CUDA(cudaSetDevice(rr.ord));
        GP_CUDA(cudaMemcpyAsync(&head, rr.scr, 16, cudaMemcpyDeviceToHost, rr.st->stream));
        GP_CUDA(cudaStreamSynchronize(rr.st->stream));
        GP_CUDA(cudaSetDevice(dd.ord));
        uint64_t land_sum = 0;
        uint32_t land_xor = 0;
        int r = run_digest(*dd.st, dd.dst, nwords, &land_sum, &land_xor, dd.st->stream, err, errlen);
        if (r) return r;
        expect[s].wait();
        c.checksum_expected = D[s].exp_sum;
        c.xor_expected = D[s].exp_xor;
        if (ce) {  // no kernel saw the words in flight: the landed copy's digest is the checksum
            head.out_sum = land_sum;
            head.out_xor = land_xor;
        }
        c.checksum_got = head.out_sum;
        c.xor_got = head.out_xor;
        c.verified = (head.out_sum == D[s].exp_sum && head.out_xor == D[s].exp_xor && land_sum == D[s].exp_sum &&
                      land_xor == D[s].exp_xor)
                         ? 1
                         : 0;
        c.ms_best = *std::min_element(ms.begin(), ms.end());
        c.ms_median = median_of(ms);
        c.gbps = c.ms_median > 0 ? bytes / (c.ms_median * 1e6) : 0;
        c.frac_of_peak = c.gbps / peak;
        return GPUPROBE_OK;
    };

    if (serial) {
        for (int s = 0; s < nd && rc == GPUPROBE_OK; ++s)
            for (int t = 0; t < nd; ++t) {
                if (s == t || !cells[s * nd + t].peer_ok) continue;
                P2PDev &dd = D[t];
                P2PDev &rr = D[runner(s, t)];
                GP_TRY(cudaSetDevice(dd.ord));
                GP_TRY(cudaMemsetAsync(dd.dst, 0, bytes, dd.st->stream));
                GP_TRY(cudaStreamSynchronize(dd.st->stream));  // PUSH writes it from the source's stream
                for (int i = 0; i < cfg->warmup; ++i) GP_TRY(launch_pair(s, t, false));
                std::vector<float> ms(cfg->iters);
                for (int i = 0; i < cfg->iters; ++i) {
                    GP_TRY(launch_pair(s, t, true));
                    GP_TRY(cudaEventSynchronize(rr.ev1));
                    GP_TRY(cudaEventElapsedTime(&ms[i], rr.ev0, rr.ev1));
                }
                rc = finish_pair(s, t, ms);
                if (rc) break;
            }
    } else {
        for (int r = 1; r < nd && rc == GPUPROBE_OK; ++r) {
            // round r: data flows s -> (s + r) mod n for every s at once, so each GPU sends and receives simultaneously
            std::vector<std::vector<float>> ms(nd);
            for (int s = 0; s < nd; ++s) {
                const int t = (s + r) % nd;
                if (!cells[s * nd + t].peer_ok) continue;
                GP_TRY(cudaSetDevice(D[t].ord));
                GP_TRY(cudaMemsetAsync(D[t].dst, 0, bytes, D[t].st->stream));
                GP_TRY(cudaStreamSynchronize(D[t].st->stream));  // PUSH writes it from another device's stream
            }
            for (int i = 0; i < cfg->warmup + cfg->iters; ++i) {
                const bool timed = i >= cfg->warmup;
                for (int s = 0; s < nd; ++s) {
                    const int t = (s + r) % nd;
                    if (cells[s * nd + t].peer_ok) GP_TRY(launch_pair(s, t, timed));
                }
                for (int s = 0; s < nd; ++s) {
                    const int t = (s + r) % nd;
                    if (!cells[s * nd + t].peer_ok) continue;
                    P2PDev &rr = D[runner(s, t)];
                    GP_TRY(cudaSetDevice(rr.ord));
                    GP_TRY(cudaStreamSynchronize(rr.st->stream));
                    if (timed) {
                        float m = 0;
                        GP_TRY(cudaEventElapsedTime(&m, rr.ev0, rr.ev1));
                        ms[s].push_back(m);
                    }
                }
            }
            for (int s = 0; s < nd; ++s) {
                const int t = (s + r) % nd;
                if (!cells[s * nd + t].peer_ok) continue;
                rc = finish_pair(s, t, ms[s]);
                if (rc) break;
            }
        }
    }
    cleanup();
    if (rc) return rc;
    bool bad = false;
    for (int s = 0; s < nd; ++s)
        for (int t = 0; t < nd; ++t)
            if (s != t && cells[s * nd + t].peer_ok && !cells[s * nd + t].verified) {
                if (!bad)
                    set_err(err, errlen, "p2p %d -> %d: checksum %016llx, expected %016llx", cells[s * nd + t].src,
                            cells[s * nd + t].dst, (unsigned long long)cells[s * nd + t].checksum_got,
                            (unsigned long long)cells[s * nd + t].checksum_expected);
                bad = true;
            }
    if (bad) return GPUPROBE_FAIL_VERIFY;
    if (missing_peer) {
        set_err(err, errlen, "at least one device pair has no peer access");
        return GPUPROBE_ENOPEER;
    }
    return GPUPROBE_OK;
#undef GP_TRY
}

// The P2P kernels with source and landing buffer on ONE device: no link is crossed, so this says nothing about NVLink --
// it exists so that the same kernels and their checksum reduction can be checked against the oracle on a box with a
// single GPU (the plugin workload's pod, a 1-GPU test machine). cell->gbps is then local HBM copy bandwidth.
extern "C" int gpuprobe_p2p_loopback(int dev, uint64_t bytes, int iters, int kernel, uint32_t seed, gpuprobe_p2p_cell *cell,
                                     char *err, size_t errlen) {
    if (!cell || bytes == 0 || (bytes & 15u) || iters < 1 || kernel < 0 || kernel > 1) {
        set_err(err, errlen, "bad p2p loopback arguments: bytes=%llu iters=%d kernel=%d", (unsigned long long)bytes, iters, kernel);
        return GPUPROBE_EINVAL;
    }
    std::memset(cell, 0, sizeof(*cell));
    DeviceState *st = nullptr;
    int rc = use_device(dev, &st, err, errlen);
    if (rc) return rc;
    std::lock_guard<std::mutex> lock(st->mu);
    const uint64_t nwords = bytes / 4, n16 = bytes / 16;
    uint32_t *src = nullptr, *dst = nullptr;
    P2PScratch *scr = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    auto cleanup = [&]() {
        cudaDeviceSynchronize();
        cudaFree(src);
        cudaFree(dst);
        cudaFree(scr);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
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
    GP_TRY(cudaMalloc(&src, bytes));
    GP_TRY(cudaMalloc(&dst, bytes));
    GP_TRY(cudaMalloc(&scr, sizeof(P2PScratch)));
    GP_TRY(cudaMemsetAsync(scr, 0, sizeof(P2PScratch), st->stream));
    GP_TRY(cudaMemsetAsync(dst, 0, bytes, st->stream));
    GP_TRY(cudaEventCreate(&ev0));
    GP_TRY(cudaEventCreate(&ev1));
    rc = launch_fill_words(*st, src, nwords, hash_base(seed, stream_p2p(dev)), st->stream, err, errlen);
    if (rc) {
        cleanup();
        return rc;
    }
    if (kernel == 0)
        GP_TRY(cudaFuncSetAttribute(p2p_bulk_copy_checksum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(P2PB_SMEM)));
    std::vector<float> ms(iters);
    for (int i = 0; i < iters + 1; ++i) {  // one untimed launch first
        if (i) GP_TRY(cudaEventRecord(ev0, st->stream));
        if (kernel == 0)
            p2p_bulk_copy_checksum_kernel<<<p2p_bulk_grid(*st, bytes), P2PB_THREADS, P2PB_SMEM, st->stream>>>(
                reinterpret_cast<const unsigned char *>(src), reinterpret_cast<unsigned char *>(dst), bytes, scr);
        else
            p2p_copy_checksum_kernel<<<p2p_grid(*st, n16), P2P_THREADS, 0, st->stream>>>(reinterpret_cast<const uint4 *>(src),
                                                                                         reinterpret_cast<uint4 *>(dst), n16, scr);
        GP_TRY(cudaGetLastError());
        if (i) {
            GP_TRY(cudaEventRecord(ev1, st->stream));
            GP_TRY(cudaEventSynchronize(ev1));
            GP_TRY(cudaEventElapsedTime(&ms[i - 1], ev0, ev1));
        }
    }
    P2PScratch head{};
    GP_TRY(cudaMemcpyAsync(&head, scr, 16, cudaMemcpyDeviceToHost, st->stream));
    GP_TRY(cudaStreamSynchronize(st->stream));
    uint64_t land_sum = 0;
    uint32_t land_xor = 0;
    rc = run_digest(*st, dst, nwords, &land_sum, &land_xor, st->stream, err, errlen);
    if (rc) {
        cleanup();
        return rc;
    }
    cell->src = cell->dst = dev;
    cell->peer_ok = 1;
    gpuprobe_p2p_expected(nwords, seed, dev, 0, &cell->checksum_expected, &cell->xor_expected);
    cell->checksum_got = head.out_sum;
    cell->xor_got = head.out_xor;
    cell->verified = (head.out_sum == cell->checksum_expected && head.out_xor == cell->xor_expected && land_sum == cell->checksum_expected &&
                      land_xor == cell->xor_expected)
                         ? 1
                         : 0;
    cell->ms_best = *std::min_element(ms.begin(), ms.end());
    cell->ms_median = median_of(ms);
    cell->gbps = cell->ms_median > 0 ? bytes / (cell->ms_median * 1e6) : 0;
    cell->frac_of_peak = 0;  // no link involved
    cleanup();
    if (!cell->verified) {
        set_err(err, errlen, "p2p loopback on device %d: checksum %016llx, expected %016llx", dev, (unsigned long long)cell->checksum_got,
                (unsigned long long)cell->checksum_expected);
        return GPUPROBE_FAIL_VERIFY;
    }
    return GPUPROBE_OK;
#undef GP_TRY
}

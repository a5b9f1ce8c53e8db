// p2p_bidir_exp.cu -- scratch experiments behind the P2P kernel's shape when BOTH directions of the links are
// loaded (not part of the product).  GPU 0 and GPU 1 each move 1 GiB to the other at the same time; every shape is
// run as PULL (kernel on the destination, loads cross NVLink) and as PUSH (kernel on the source, stores cross NVLink):
//   ldst   : register path, 128- or 256-bit accesses, U vectors in flight per thread
//   bulk   : cp.async.bulk global -> shared -> global through an mbarrier ring (the copy engine inside the SM issues the
//            transfers: CHUNK bytes per request instead of one 128-byte line per warp instruction)
//   split  : half of the bytes pulled by the destination while the source pushes the other half
// One direction alone is measured too, as the reference point.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/exp/bin/p2p_bidir_exp tools/exp/p2p_bidir_exp.cu
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__); exit(2); } } while (0)

template <int VB> struct V;
template <> struct V<16> { uint32_t w[4]; };
template <> struct V<32> { uint32_t w[8]; };

template <int VB> __device__ __forceinline__ V<VB> ldv(const void *p) {
    V<VB> r;
    if constexpr (VB == 16)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    else
        asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]) : "l"(p));
    return r;
}
template <int VB> __device__ __forceinline__ void stv(void *p, const V<VB> &r) {
    if constexpr (VB == 16)
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
    else
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]), "r"(r.w[4]), "r"(r.w[5]), "r"(r.w[6]), "r"(r.w[7]) : "memory");
}

template <int VB, int U, int T>
__global__ void __launch_bounds__(T) k_ldst(const char *__restrict__ src, char *__restrict__ dst, size_t bytes, unsigned long long *sum) {
    unsigned long long s = 0;
    const size_t nvec = bytes / VB, tile = (size_t)T * U, n_tiles = nvec / tile;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        V<VB> v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldv<VB>(src + (base + (size_t)u * T) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            stv<VB>(dst + (base + (size_t)u * T) * VB, v[u]);
#pragma unroll
            for (int i = 0; i < VB / 4; ++i) s += v[u].w[i];
        }
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(sum, s);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Bulk-copy ring: CHUNK-byte requests global -> shared (mbarrier complete_tx), all threads checksum the stage out of
// shared memory, one thread sends it on with a bulk store. A stage is refilled one iteration late, after
// wait_group.read has shown that the store that read it has drained.
template <int CHUNK, int STAGES, int T, bool SUM>
__global__ void __launch_bounds__(T) k_bulk(const char *__restrict__ src, char *__restrict__ dst, size_t bytes, unsigned long long *sum) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)STAGES * CHUNK);
    const size_t n_chunks = bytes / CHUNK;
    const size_t first = blockIdx.x, stride = gridDim.x;
    const size_t mine = first < n_chunks ? (n_chunks - first + stride - 1) / stride : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto load = [&](size_t k) {
        const int s = (int)(k % STAGES);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem + (size_t)s * CHUNK)),
                     "l"(src + (first + k * stride) * CHUNK), "r"(CHUNK), "r"(smem_u32(&full[s])) : "memory");
    };
    if (!SUM && threadIdx.x != 0) return;  // nothing to checksum: the issuing thread alone walks the ring
    if (threadIdx.x == 0)
        for (size_t k = 0; k < (size_t)(STAGES - 1) && k < mine; ++k) load(k);
    unsigned long long acc = 0;
    for (size_t k = 0; k < mine; ++k) {
        const int s = (int)(k % STAGES);
        const uint32_t parity = (uint32_t)((k / STAGES) & 1u);
        asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W_%=;\n}\n" ::"r"(smem_u32(&full[s])), "r"(parity) : "memory");
        if (SUM) {
            const uint4 *p = reinterpret_cast<const uint4 *>(smem + (size_t)s * CHUNK);
            for (int i = threadIdx.x; i < CHUNK / 16; i += T) {
                const uint4 v = p[i];
                acc += (unsigned long long)v.x + v.y + (unsigned long long)v.z + v.w;
            }
            __syncthreads();  // every thread has read the stage before it is handed to the store
        }
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (first + k * stride) * CHUNK), "r"(smem_u32(smem + (size_t)s * CHUNK)), "r"(CHUNK) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (k + STAGES - 1 < mine) {
                // the stage about to be refilled was stored in iteration k-1: at most this iteration's store may still be reading
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                load(k + STAGES - 1);
            }
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (SUM && (threadIdx.x & 31) == 0 && acc) atomicAdd(sum, acc);
}

struct Side {
    int dev;
    char *src, *dst;  // src: this GPU's outgoing data; dst: landing buffer for the peer's data
    unsigned long long *sum;
    cudaStream_t st, st2;
    cudaEvent_t e0, e1, f0, f1;
    int sms;
};

int main(int argc, char **argv) {
    const size_t bytes = (size_t)1 << 30;
    const int iters = argc > 1 ? atoi(argv[1]) : 5;
    int n;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    Side S[2];
    for (int d = 0; d < 2; ++d) {
        S[d].dev = d;
        CK(cudaSetDevice(d));
        CK(cudaDeviceEnablePeerAccess(1 - d, 0));
        CK(cudaMalloc(&S[d].src, bytes));
        CK(cudaMalloc(&S[d].dst, bytes));
        CK(cudaMalloc(&S[d].sum, 8));
        CK(cudaMemset(S[d].src, 0x5a + d, bytes));
        CK(cudaMemset(S[d].dst, 0, bytes));
        CK(cudaStreamCreateWithFlags(&S[d].st, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&S[d].st2, cudaStreamNonBlocking));
        CK(cudaEventCreate(&S[d].e0)); CK(cudaEventCreate(&S[d].e1)); CK(cudaEventCreate(&S[d].f0)); CK(cudaEventCreate(&S[d].f1));
        CK(cudaDeviceGetAttribute(&S[d].sms, cudaDevAttrMultiProcessorCount, d));
        CK(cudaDeviceSynchronize());
    }
    // mode: 0 pull (kernel on the destination), 1 push (kernel on the source), 2 split (both, half each)
    // dirs: 1 = only 0 -> 1; 2 = both directions at once
    auto run = [&](const char *name, auto launch, int mode, int dirs) {
        std::vector<float> ms;
        for (int it = 0; it < 2 + iters; ++it) {
            for (int d = 0; d < dirs; ++d) {
                // direction d: data flows from GPU d to GPU 1-d
                Side &from = S[d], &to = S[1 - d];
                if (mode == 0 || mode == 2) {
                    const size_t part = mode == 2 ? bytes / 2 : bytes;
                    CK(cudaSetDevice(to.dev));
                    CK(cudaEventRecord(to.e0, to.st));
                    launch(to, from.src, to.dst, part, to.st);
                    CK(cudaEventRecord(to.e1, to.st));
                }
                if (mode == 1 || mode == 2) {
                    const size_t off = mode == 2 ? bytes / 2 : 0;
                    CK(cudaSetDevice(from.dev));
                    CK(cudaEventRecord(from.f0, from.st2));
                    launch(from, from.src + off, to.dst + off, bytes - off, from.st2);
                    CK(cudaEventRecord(from.f1, from.st2));
                }
            }
            float worst = 0;
            for (int d = 0; d < 2; ++d) {
                CK(cudaSetDevice(d));
                CK(cudaDeviceSynchronize());
            }
            for (int d = 0; d < dirs; ++d) {
                Side &from = S[d], &to = S[1 - d];
                float m = 0, m2 = 0;
                if (mode == 0 || mode == 2) CK(cudaEventElapsedTime(&m, to.e0, to.e1));
                if (mode == 1 || mode == 2) CK(cudaEventElapsedTime(&m2, from.f0, from.f1));
                worst = std::max(worst, std::max(m, m2));
            }
            if (it >= 2) ms.push_back(worst);
        }
        std::sort(ms.begin(), ms.end());
        printf("{\"exp\": \"%s\", \"mode\": \"%s\", \"dirs\": %d, \"gbps_per_dir_median\": %.1f, \"gbps_per_dir_best\": %.1f}\n", name,
               mode == 0 ? "pull" : mode == 1 ? "push" : "split", dirs, bytes / (ms[ms.size() / 2] * 1e6), bytes / (ms[0] * 1e6));
        fflush(stdout);
    };
#define LDST(VB, U, T, CT) [&](Side &r, const char *s, char *d, size_t b, cudaStream_t st) { k_ldst<VB, U, T><<<r.sms * CT, T, 0, st>>>(s, d, b, r.sum); CK(cudaGetLastError()); }
#define BULK(CH, SG, T, CT, SUM) [&](Side &r, const char *s, char *d, size_t b, cudaStream_t st) { \
        const size_t sm = (size_t)SG * CH + SG * 8; \
        CK(cudaFuncSetAttribute(k_bulk<CH, SG, T, SUM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
        k_bulk<CH, SG, T, SUM><<<r.sms * CT, T, sm, st>>>(s, d, b, r.sum); CK(cudaGetLastError()); }

    for (int dirs : {1, 2}) {
        if (getenv("P2P_EXP_CE_ONLY")) break;
        for (int mode : {0, 1}) {
            run("ldst_v16_u8_t512_x4", LDST(16, 8, 512, 4), mode, dirs);   // the product kernel's shape
            run("ldst_v32_u4_t512_x4", LDST(32, 4, 512, 4), mode, dirs);
            run("ldst_v32_u8_t512_x2", LDST(32, 8, 512, 2), mode, dirs);
            run("ldst_v16_u4_t512_x2", LDST(16, 4, 512, 2), mode, dirs);
            run("ldst_v16_u16_t512_x2", LDST(16, 16, 512, 2), mode, dirs);
            run("bulk_4k_s8_x4", BULK(4096, 8, 128, 4, true), mode, dirs);
            run("bulk_8k_s6_x4", BULK(8192, 6, 128, 4, true), mode, dirs);
            run("bulk_16k_s6_x2", BULK(16384, 6, 128, 2, true), mode, dirs);
            run("bulk_32k_s6_x1", BULK(32768, 6, 128, 1, true), mode, dirs);
            run("bulk_32k_s3_x2", BULK(32768, 3, 128, 2, true), mode, dirs);
            run("bulk_16k_s12_x1", BULK(16384, 12, 128, 1, true), mode, dirs);
            run("bulk_16k_s6_x2_nosum", BULK(16384, 6, 128, 2, false), mode, dirs);
            run("bulk_2k_s16_x4", BULK(2048, 16, 128, 4, true), mode, dirs);
            run("bulk_1k_s16_x8", BULK(1024, 16, 128, 8, true), mode, dirs);
        }
    }
    // the copy engines, for reference: what the hardware moves without any SM in the path (no checksum possible)
    for (int dirs : {1, 2}) {
        run("copy_engine_memcpy", [&](Side &, const char *s, char *d, size_t b, cudaStream_t st) { CK(cudaMemcpyAsync(d, s, b, cudaMemcpyDeviceToDevice, st)); }, 0, dirs);
        run("copy_engine_memcpy", [&](Side &, const char *s, char *d, size_t b, cudaStream_t st) { CK(cudaMemcpyAsync(d, s, b, cudaMemcpyDeviceToDevice, st)); }, 1, dirs);
    }
    if (getenv("P2P_EXP_CE_ONLY")) return 0;
    run("split_ldst_v16_u8_t512_x2", LDST(16, 8, 512, 2), 2, 2);
    run("split_bulk_16k_s6_x1", BULK(16384, 6, 128, 1, true), 2, 2);
    return 0;
}

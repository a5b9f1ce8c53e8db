// vadd_exp.cu -- scratch experiments behind the vectorAdd kernel choices (not part of the product).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --fmad=false -o tools/exp/bin/vadd_exp tools/exp/vadd_exp.cu
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(2); } } while (0)

enum LoadHint { L_NC_NA = 0, L_PLAIN = 1, L_CS = 2, L_LU = 3, L_NC = 4, L_NC_NA_PF256 = 5, L_CG = 6 };
enum StoreHint { S_NA = 0, S_PLAIN = 1, S_CS = 2, S_WT = 3, S_CG = 4 };

template <int LH> __device__ __forceinline__ float4 ld4(const float *p) {
    float4 r;
    if constexpr (LH == L_NC_NA) asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else if constexpr (LH == L_PLAIN) asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else if constexpr (LH == L_CS) asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else if constexpr (LH == L_LU) asm volatile("ld.global.lu.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else if constexpr (LH == L_NC) asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else if constexpr (LH == L_NC_NA_PF256) asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    else asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
template <int SH> __device__ __forceinline__ void st4(float *p, float4 v) {
    if constexpr (SH == S_NA) asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else if constexpr (SH == S_PLAIN) asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else if constexpr (SH == S_CS) asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else if constexpr (SH == S_WT) asm volatile("st.global.wt.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else asm volatile("st.global.cg.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 add4(float4 x, float4 y) { return make_float4(__fadd_rn(x.x, y.x), __fadd_rn(x.y, y.y), __fadd_rn(x.z, y.z), __fadd_rn(x.w, y.w)); }

// generic tile kernel: T threads, U float4 per input per thread
template <int T, int U, int LH, int SH>
__global__ void __launch_bounds__(T) k_vadd(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * U * 4) + threadIdx.x * 4;
    float4 va[U], vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) va[u] = ld4<LH>(a + base + (size_t)u * T * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) vb[u] = ld4<LH>(b + base + (size_t)u * T * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) st4<SH>(c + base + (size_t)u * T * 4, add4(va[u], vb[u]));
}
// 256-bit
template <int T, int U>
__global__ void __launch_bounds__(T) k_vadd8(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * U * 8) + threadIdx.x * 8;
    float va[U][8], vb[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(va[u][0]), "=f"(va[u][1]), "=f"(va[u][2]), "=f"(va[u][3]), "=f"(va[u][4]), "=f"(va[u][5]), "=f"(va[u][6]), "=f"(va[u][7]) : "l"(a + base + (size_t)u * T * 8));
#pragma unroll
    for (int u = 0; u < U; ++u) asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(vb[u][0]), "=f"(vb[u][1]), "=f"(vb[u][2]), "=f"(vb[u][3]), "=f"(vb[u][4]), "=f"(vb[u][5]), "=f"(vb[u][6]), "=f"(vb[u][7]) : "l"(b + base + (size_t)u * T * 8));
#pragma unroll
    for (int u = 0; u < U; ++u) {
        float r[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = __fadd_rn(va[u][i], vb[u][i]);
        asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(c + base + (size_t)u * T * 8), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7]) : "memory");
    }
}
// 256-bit with explicit L2 eviction priorities: LP/SP in {0: none, 1: evict_first, 2: evict_last, 3: evict_normal}
#define LD8(HINT) asm volatile("ld.global.nc.L1::no_allocate" HINT ".v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p))
#define ST8(HINT) asm volatile("st.global.L1::no_allocate" HINT ".v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory")
template <int LP> __device__ __forceinline__ void ld8p(float *v, const float *p) {
    if constexpr (LP == 1) LD8(".L2::evict_first"); else if constexpr (LP == 2) LD8(".L2::evict_last"); else if constexpr (LP == 3) LD8(".L2::evict_normal"); else LD8("");
}
template <int SP> __device__ __forceinline__ void st8p(float *p, const float *v) {
    if constexpr (SP == 1) ST8(".L2::evict_first"); else if constexpr (SP == 2) ST8(".L2::evict_last"); else if constexpr (SP == 3) ST8(".L2::evict_normal"); else ST8("");
}
template <int T, int LP, int SP>
__global__ void __launch_bounds__(T) k_vadd8_pol(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * 8) + threadIdx.x * 8;
    float x[8], y[8], r[8];
    ld8p<LP>(x, a + base);
    ld8p<LP>(y, b + base);
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __fadd_rn(x[i], y[i]);
    st8p<SP>(c + base, r);
}

// vadd with an L2 bulk prefetch of the tile AHEAD tiles further on (one thread per CTA issues it)
template <int T>
__global__ void __launch_bounds__(T) k_vadd_pf(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, size_t n_tiles, size_t ahead) {
    const size_t tile = blockIdx.x;
    if (threadIdx.x == 0 && tile + ahead < n_tiles) {
        const float *pa = a + (tile + ahead) * (T * 4), *pb = b + (tile + ahead) * (T * 4);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(pa), "r"(T * 16) : "memory");
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(pb), "r"(T * 16) : "memory");
    }
    const size_t base = tile * (T * 4) + threadIdx.x * 4;
    float4 x = ld4<L_NC_NA>(a + base), y = ld4<L_NC_NA>(b + base);
    st4<S_NA>(c + base, add4(x, y));
}

// read-only (2 streams), write-only, copy
template <int T, int U>
__global__ void __launch_bounds__(T) k_read2(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * U * 4) + threadIdx.x * 4;
    float4 acc = make_float4(0, 0, 0, 0);
#pragma unroll
    for (int u = 0; u < U; ++u) { float4 x = ld4<L_NC_NA>(a + base + (size_t)u * T * 4), y = ld4<L_NC_NA>(b + base + (size_t)u * T * 4); acc = add4(acc, add4(x, y)); }
    if (acc.x == 123.456f && acc.y == 1e30f) st4<S_NA>(c + base, acc);
}
template <int T, int U>
__global__ void __launch_bounds__(T) k_write(const float *__restrict__, const float *__restrict__, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * U * 4) + threadIdx.x * 4;
#pragma unroll
    for (int u = 0; u < U; ++u) st4<S_NA>(c + base + (size_t)u * T * 4, make_float4(1.f, 2.f, 3.f, (float)threadIdx.x));
}
template <int T, int U>
__global__ void __launch_bounds__(T) k_copy(const float *__restrict__ a, const float *__restrict__, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * U * 4) + threadIdx.x * 4;
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld4<L_NC_NA>(a + base + (size_t)u * T * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) st4<S_NA>(c + base + (size_t)u * T * 4, v[u]);
}
// persistent, chunked: CTA j owns a contiguous run of tiles
template <int T, int U>
__global__ void __launch_bounds__(T) k_vadd_chunked(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, size_t n_tiles) {
    const size_t per = (n_tiles + gridDim.x - 1) / gridDim.x;
    const size_t lo = per * blockIdx.x, hi = min(lo + per, n_tiles);
    for (size_t t = lo; t < hi; ++t) {
        const size_t base = t * (T * U * 4) + threadIdx.x * 4;
        float4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld4<L_NC_NA>(a + base + (size_t)u * T * 4);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld4<L_NC_NA>(b + base + (size_t)u * T * 4);
#pragma unroll
        for (int u = 0; u < U; ++u) st4<S_NA>(c + base + (size_t)u * T * 4, add4(va[u], vb[u]));
    }
}

// ---- TMA (cp.async.bulk) experiments -----------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mb_expect(uint64_t *b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mb_wait(uint64_t *b, uint32_t ph) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W_%=;\n}\n" :: "r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void g2s(void *sm, const void *g, uint32_t n, uint64_t *b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(s32(sm)), "l"(g), "r"(n), "r"(s32(b)) : "memory");
}
__device__ __forceinline__ void s2g(void *g, const void *sm, uint32_t n) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(g), "r"(s32(sm)), "r"(n) : "memory");
}
// pure DMA copy: one thread per CTA moves TILE_B-byte tiles global -> shared -> global through a ring
template <int TILE_B, int STAGES>
__global__ void __launch_bounds__(32) k_tma_copy(const float *__restrict__ a, const float *__restrict__, float *__restrict__ c, size_t n_tiles) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t *full = (uint64_t *)(sm + (size_t)STAGES * TILE_B);
    if (threadIdx.x != 0) return;
    for (int s = 0; s < STAGES; ++s) mb_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t first = blockIdx.x, stride = gridDim.x;
    const size_t mine = first < n_tiles ? (n_tiles - first + stride - 1) / stride : 0;
    const char *src = (const char *)a; char *dst = (char *)c;
    for (size_t k = 0; k < STAGES && k < mine; ++k) { mb_expect(&full[k], TILE_B); g2s(sm + k * TILE_B, src + (first + k * stride) * TILE_B, TILE_B, &full[k]); }
    for (size_t k = 0; k < mine; ++k) {
        const int s = k % STAGES;
        mb_wait(&full[s], (k / STAGES) & 1);
        s2g(dst + (first + k * stride) * TILE_B, sm + (size_t)s * TILE_B, TILE_B);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        if (k + STAGES < mine) {
            // the store just issued must have finished READING stage s before it is refilled
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            mb_expect(&full[s], TILE_B);
            g2s(sm + (size_t)s * TILE_B, src + (first + (k + STAGES) * stride) * TILE_B, TILE_B, &full[s]);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// vadd: LDG loads (many CTAs in flight), result staged in shared memory, one bulk store per CTA tile
template <int T, int U>
__global__ void __launch_bounds__(T) k_vadd_bulkstore(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    __shared__ __align__(128) float4 tile[T * U];
    const size_t base = (size_t)blockIdx.x * (T * U * 4) + threadIdx.x * 4;
    float4 va[U], vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) va[u] = ld4<L_NC_NA>(a + base + (size_t)u * T * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) vb[u] = ld4<L_NC_NA>(b + base + (size_t)u * T * 4);
#pragma unroll
    for (int u = 0; u < U; ++u) tile[u * T + threadIdx.x] = add4(va[u], vb[u]);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        s2g(c + (size_t)blockIdx.x * (T * U * 4), tile, T * U * 16);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}


// ---- round 2: DRAM read/write turnaround experiments -----------------------------------------------------
__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Chip-wide temporal separation of reads and writes: persistent CTAs, each round pulls STAGES tiles of A and of B into
// shared memory with bulk loads, adds in place, and sends the STAGES result tiles back with bulk stores. With period_ns > 0
// the issuing thread only starts loads while (globaltimer % period_ns) < read_ns and only starts stores in the rest of
// the period, so every SM of the chip reads in the same window and writes in the same window. period_ns = 0: same
// structure, no windows (the control).
template <int TILE_F, int STAGES, int T>
__global__ void __launch_bounds__(T) k_vadd_phased(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, size_t n_tiles,
                                                   unsigned period_ns, unsigned read_ns) {
    extern __shared__ __align__(128) unsigned char sm[];
    constexpr uint32_t TILE_B = TILE_F * 4u;
    float *sa = (float *)sm, *sb = (float *)(sm + (size_t)STAGES * TILE_B);
    uint64_t *full = (uint64_t *)(sm + 2 * (size_t)STAGES * TILE_B);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) mb_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t n_groups = n_tiles / STAGES;  // n_tiles is a multiple of STAGES in this harness
    uint32_t round = 0;
    for (size_t g = blockIdx.x; g < n_groups; g += gridDim.x, ++round) {
        const size_t t0 = g * STAGES;
        if (threadIdx.x == 0) {
            if (period_ns) while ((unsigned)(gtimer_ns() % period_ns) >= read_ns) {}
            for (int s = 0; s < STAGES; ++s) {
                mb_expect(&full[s], 2 * TILE_B);
                g2s(sa + (size_t)s * TILE_F, a + (t0 + s) * TILE_F, TILE_B, &full[s]);
                g2s(sb + (size_t)s * TILE_F, b + (t0 + s) * TILE_F, TILE_B, &full[s]);
            }
        }
        for (int s = 0; s < STAGES; ++s) {
            mb_wait(&full[s], round & 1);
            float4 *pa = (float4 *)(sa + (size_t)s * TILE_F);
            const float4 *pb = (const float4 *)(sb + (size_t)s * TILE_F);
#pragma unroll
            for (int i = threadIdx.x; i < TILE_F / 4; i += T) pa[i] = add4(pa[i], pb[i]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            if (period_ns) while ((unsigned)(gtimer_ns() % period_ns) < read_ns) {}
            for (int s = 0; s < STAGES; ++s) s2g(c + (t0 + s) * TILE_F, sa + (size_t)s * TILE_F, TILE_B);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// Tile order: does DRAM page locality between co-resident CTAs matter? CTA i handles tile perm(i), where perm
// bit-reverses the low WBITS bits of i (a scatter inside windows of 2^WBITS tiles); WBITS = 0 is the identity.
template <int T, int WBITS>
__global__ void __launch_bounds__(T) k_vadd_swz(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    unsigned i = blockIdx.x;
    if constexpr (WBITS > 0) {
        const unsigned low = i & ((1u << WBITS) - 1u);
        i = (i & ~((1u << WBITS) - 1u)) | (__brev(low) >> (32 - WBITS));
    }
    const size_t base = (size_t)i * (T * 4) + threadIdx.x * 4;
    float4 x = ld4<L_NC_NA>(a + base), y = ld4<L_NC_NA>(b + base);
    st4<S_NA>(c + base, add4(x, y));
}
// Same 12 B/element, but each CTA first reads a whole run of RUN tiles (A and B back to back, registers), then
// writes the RUN results: per-CTA separation of a read burst from a write burst, without shared memory.
template <int T, int RUN>
__global__ void __launch_bounds__(T) k_vadd_burst(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    const size_t base = (size_t)blockIdx.x * (T * 4 * RUN) + threadIdx.x * 4;
    float4 x[RUN], y[RUN];
#pragma unroll
    for (int r = 0; r < RUN; ++r) x[r] = ld4<L_NC_NA>(a + base + (size_t)r * T * 4);
#pragma unroll
    for (int r = 0; r < RUN; ++r) y[r] = ld4<L_NC_NA>(b + base + (size_t)r * T * 4);
#pragma unroll
    for (int r = 0; r < RUN; ++r) x[r] = add4(x[r], y[r]);
#pragma unroll
    for (int r = 0; r < RUN; ++r) st4<S_NA>(c + base + (size_t)r * T * 4, x[r]);
}

struct Exp { std::string name; double bytes_per_elem; std::function<void(const float *, const float *, float *, size_t, cudaStream_t)> run; };

template <int T, int U, int LH, int SH> Exp mk(const char *name) {
    return {name, 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd<T, U, LH, SH><<<(unsigned)(n / (T * U * 4)), T, 0, s>>>(a, b, c); }};
}
template <int T, int U> Exp mk8(const char *name) {
    return {name, 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd8<T, U><<<(unsigned)(n / (T * U * 8)), T, 0, s>>>(a, b, c); }};
}

int main(int argc, char **argv) {
    int lg = argc > 1 ? atoi(argv[1]) : 30;
    int iters = argc > 2 ? atoi(argv[2]) : 20;
    const char *filter = argc > 3 ? argv[3] : "";
    const size_t n = (size_t)1 << lg;
    if (const char *g = getenv("EXP_L2_FETCH")) { CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(g))); size_t v = 0; cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity); fprintf(stderr, "L2 fetch granularity now %zu\n", v); }
    const size_t PAD = 64u << 20;  // room for offset experiments
    char *pool;
    CK(cudaMalloc(&pool, 3 * (n * 4 + PAD)));
    CK(cudaMemset(pool, 0, 3 * (n * 4 + PAD)));
    cudaStream_t s; CK(cudaStreamCreate(&s));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));

    std::vector<Exp> E;
    E.push_back({"read2_t256_u2", 8, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_read2<256, 2><<<(unsigned)(n / (256 * 2 * 4)), 256, 0, s>>>(a, b, c); }});
    E.push_back({"write_t256_u2", 4, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_write<256, 2><<<(unsigned)(n / (256 * 2 * 4)), 256, 0, s>>>(a, b, c); }});
    E.push_back({"copy_t256_u2", 8, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_copy<256, 2><<<(unsigned)(n / (256 * 2 * 4)), 256, 0, s>>>(a, b, c); }});
    E.push_back(mk<256, 2, L_NC_NA, S_NA>("vadd_t256_u2"));
    E.push_back(mk<256, 1, L_NC_NA, S_NA>("vadd_t256_u1"));
    E.push_back(mk<512, 1, L_NC_NA, S_NA>("vadd_t512_u1"));
    E.push_back(mk<1024, 1, L_NC_NA, S_NA>("vadd_t1024_u1"));
    E.push_back(mk<128, 1, L_NC_NA, S_NA>("vadd_t128_u1"));
    E.push_back(mk<128, 2, L_NC_NA, S_NA>("vadd_t128_u2"));
    E.push_back(mk<512, 2, L_NC_NA, S_NA>("vadd_t512_u2"));
    E.push_back(mk<64, 2, L_NC_NA, S_NA>("vadd_t64_u2"));
    E.push_back(mk<64, 4, L_NC_NA, S_NA>("vadd_t64_u4"));
    E.push_back(mk8<128, 1>("vadd8_t128_u1"));
    E.push_back(mk8<256, 1>("vadd8_t256_u1"));
    E.push_back(mk8<512, 1>("vadd8_t512_u1"));
    E.push_back(mk<256, 2, L_PLAIN, S_NA>("vadd_ldplain"));
    E.push_back(mk<256, 2, L_CS, S_NA>("vadd_ldcs"));
    E.push_back(mk<256, 2, L_LU, S_NA>("vadd_ldlu"));
    E.push_back(mk<256, 2, L_NC, S_NA>("vadd_ldnc"));
    E.push_back(mk<256, 2, L_CG, S_NA>("vadd_ldcg"));
    E.push_back(mk<256, 2, L_NC_NA_PF256, S_NA>("vadd_ldpf256"));
    E.push_back(mk<256, 2, L_NC_NA, S_PLAIN>("vadd_stplain"));
    E.push_back(mk<256, 2, L_NC_NA, S_CS>("vadd_stcs"));
    E.push_back(mk<256, 2, L_NC_NA, S_WT>("vadd_stwt"));
    E.push_back(mk<256, 2, L_NC_NA, S_CG>("vadd_stcg"));
    E.push_back(mk<256, 2, L_CS, S_CS>("vadd_ldcs_stcs"));
#define POL(LP, SP, NAME) E.push_back({NAME, 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd8_pol<128, LP, SP><<<(unsigned)(n / (128 * 8)), 128, 0, s>>>(a, b, c); }})
    POL(0, 0, "pol_ld-none_st-none");
    POL(1, 0, "pol_ld-first_st-none");
    POL(1, 2, "pol_ld-first_st-last");
    POL(0, 2, "pol_ld-none_st-last");
    POL(1, 1, "pol_ld-first_st-first");
    POL(2, 1, "pol_ld-last_st-first");
    POL(3, 3, "pol_ld-normal_st-normal");
    POL(0, 1, "pol_ld-none_st-first");
    for (size_t mb : {1, 4, 16, 64, 256}) {
        const size_t ahead = mb * 1024 * 1024 / 4096;  // tiles of 4 KiB per array
        E.push_back({"vadd_pfL2_" + std::to_string(mb) + "MB", 12, [ahead](const float *a, const float *b, float *c, size_t n, cudaStream_t s) {
            k_vadd_pf<256><<<(unsigned)(n / 1024), 256, 0, s>>>(a, b, c, n / 1024, ahead); }});
    }
    for (int cl : {2, 4, 8}) {
        E.push_back({"vadd_cluster" + std::to_string(cl), 12, [cl](const float *a, const float *b, float *c, size_t n, cudaStream_t s) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)(n / 1024)); cfg.blockDim = dim3(256); cfg.stream = s;
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, k_vadd<256, 1, L_NC_NA, S_NA>, a, b, c); }});
    }
    E.push_back({"vadd_bulkstore_t256_u1", 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_bulkstore<256, 1><<<(unsigned)(n / (256 * 1 * 4)), 256, 0, s>>>(a, b, c); }});
    E.push_back({"vadd_bulkstore_t256_u2", 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_bulkstore<256, 2><<<(unsigned)(n / (256 * 2 * 4)), 256, 0, s>>>(a, b, c); }});
    E.push_back({"vadd_bulkstore_t512_u2", 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_bulkstore<512, 2><<<(unsigned)(n / (512 * 2 * 4)), 512, 0, s>>>(a, b, c); }});
    {
        auto add_tma = [&](const char *name, auto kern, int tile_b, int stages, int ctas) {
            size_t smem = (size_t)stages * tile_b + stages * 8;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            E.push_back({name, 8, [=](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { kern<<<sms * ctas, 32, smem, s>>>(a, b, c, n * 4 / tile_b); }});
        };
        add_tma("tmacopy_16k_s8_x1", k_tma_copy<16384, 8>, 16384, 8, 1);
        add_tma("tmacopy_16k_s4_x3", k_tma_copy<16384, 4>, 16384, 4, 3);
        add_tma("tmacopy_32k_s6_x1", k_tma_copy<32768, 6>, 32768, 6, 1);
        add_tma("tmacopy_8k_s8_x3", k_tma_copy<8192, 8>, 8192, 8, 3);
        add_tma("tmacopy_4k_s8_x6", k_tma_copy<4096, 8>, 4096, 8, 6);
    }
    for (int k : {1, 2, 4, 8, 16})
        E.push_back({"chunked_x" + std::to_string(k), 12, [k, sms](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_chunked<256, 2><<<sms * k, 256, 0, s>>>(a, b, c, n / 2048); }});

    // round 2: read/write phase separation and tile order
    {
        auto add_phased = [&](const char *name, auto kern, int tile_f, int stages, int ctas, unsigned period, unsigned readns) {
            size_t smem = 2 * (size_t)stages * tile_f * 4 + stages * 8;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            E.push_back({name, 12, [=](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { kern<<<sms * ctas, 256, smem, s>>>(a, b, c, n / tile_f, period, readns); }});
        };
        add_phased("phased_16k_s4_x1_nowin", k_vadd_phased<4096, 4, 256>, 4096, 4, 1, 0, 0);
        add_phased("phased_16k_s4_x1_p4us", k_vadd_phased<4096, 4, 256>, 4096, 4, 1, 4096, 2731);
        add_phased("phased_16k_s4_x1_p8us", k_vadd_phased<4096, 4, 256>, 4096, 4, 1, 8192, 5461);
        add_phased("phased_16k_s4_x1_p16us", k_vadd_phased<4096, 4, 256>, 4096, 4, 1, 16384, 10923);
        add_phased("phased_16k_s6_x1_nowin", k_vadd_phased<4096, 6, 256>, 4096, 6, 1, 0, 0);
        add_phased("phased_16k_s6_x1_p8us", k_vadd_phased<4096, 6, 256>, 4096, 6, 1, 8192, 5461);
        add_phased("phased_8k_s4_x3_nowin", k_vadd_phased<2048, 4, 256>, 2048, 4, 3, 0, 0);
        add_phased("phased_8k_s4_x3_p4us", k_vadd_phased<2048, 4, 256>, 2048, 4, 3, 4096, 2731);
        add_phased("phased_8k_s4_x3_p8us", k_vadd_phased<2048, 4, 256>, 2048, 4, 3, 8192, 5461);
        add_phased("phased_8k_s4_x3_p32us", k_vadd_phased<2048, 4, 256>, 2048, 4, 3, 32768, 21845);
        add_phased("phased_4k_s4_x6_nowin", k_vadd_phased<1024, 4, 256>, 1024, 4, 6, 0, 0);
        add_phased("phased_4k_s4_x6_p8us", k_vadd_phased<1024, 4, 256>, 1024, 4, 6, 8192, 5461);
    }
#define SWZ(W) E.push_back({"swz_brev" #W, 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_swz<256, W><<<(unsigned)(n / 1024), 256, 0, s>>>(a, b, c); }})
    SWZ(0); SWZ(4); SWZ(8); SWZ(12); SWZ(16); SWZ(20);
#define BURST(R) E.push_back({"burst_run" #R, 12, [](const float *a, const float *b, float *c, size_t n, cudaStream_t s) { k_vadd_burst<256, R><<<(unsigned)(n / (1024 * R)), 256, 0, s>>>(a, b, c); }})
    BURST(1); BURST(2); BURST(4); BURST(8); BURST(16);

    // base-pointer offsets between the three arrays (bytes)
    struct Off { const char *name; size_t a, b, c; };
    std::vector<Off> offs = {{"off0", 0, 0, 0}};
    if (std::string(filter) == "offsets" || std::string(filter) == "all")
        offs = {{"off0", 0, 0, 0}, {"off_4k", 0, 4096, 8192}, {"off_64k", 0, 65536, 131072}, {"off_1m", 0, 1u << 20, 2u << 20},
                {"off_1m4k", 0, (1u << 20) + 4096, (2u << 20) + 8192}, {"off_odd", 0, 7u * 4096 + 512, 13u * 65536 + 1024}, {"off_32m", 0, 32u << 20, 16u << 20}};

    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (auto &o : offs) {
        float *a = (float *)(pool + o.a), *b = (float *)(pool + (n * 4 + PAD) + o.b), *c = (float *)(pool + 2 * (n * 4 + PAD) + o.c);
        for (auto &x : E) {
            if (*filter && std::string(filter) != "all" && std::string(filter) != "offsets" && x.name.find(filter) == std::string::npos) continue;
            if (std::string(filter) == "offsets" && x.name != "vadd_t256_u2") continue;
            for (int i = 0; i < 3; ++i) x.run(a, b, c, n, s);
            CK(cudaGetLastError());
            std::vector<float> ms(iters);
            for (int i = 0; i < iters; ++i) {
                CK(cudaEventRecord(e0, s));
                x.run(a, b, c, n, s);
                CK(cudaEventRecord(e1, s));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms[i], e0, e1));
            }
            std::sort(ms.begin(), ms.end());
            double med = ms[iters / 2], best = ms[0];
            printf("{\"exp\": \"%s\", \"offs\": \"%s\", \"log2n\": %d, \"ms_median\": %.5f, \"gbps_median\": %.1f, \"gbps_best\": %.1f}\n", x.name.c_str(), o.name, lg, med,
                   x.bytes_per_elem * n / (med * 1e6), x.bytes_per_elem * n / (best * 1e6));
            fflush(stdout);
        }
    }
    return 0;
}

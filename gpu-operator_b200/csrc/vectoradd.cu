// vectoradd.cu -- probe (a): C = A + B over fp32 arrays in HBM, hand-written for sm_100a.
//
// Replaces the external `vectorAdd` sample binary the validator's workload pods run
// (reference validator/cuda-workload-validation.yaml:19-20, validator/plugin-workload-validation.yaml:19-20;
// installed by validator/Dockerfile:33-35,49).  The sample's kernel is 196 blocks x 256 threads of scalar
// 4-byte loads over 50,000 elements; this one is sized for the whole chip:
//
//   * HBM-bound, 12 algorithmic bytes per element (2 x 4 B read + 4 B write), 1 flop -> no tensor cores.
//   * tile kernels: every CTA owns one contiguous tile, every thread keeps UNROLL 128-bit (or 256-bit,
//     sm_100 only) loads of A and of B in flight before the first add, stores are full 128-byte lines,
//     loads bypass L1 (read once), no bounds checks on the fast path (the ragged tail is a second launch).
//   * a cp.async.bulk (TMA engine) variant that streams tiles through a shared-memory ring with
//     mbarrier completion, persistent CTAs (one or two per SM).
//   * fp32 add is __fadd_rn: IEEE round-to-nearest-even, never fused, denormals kept (no -use_fast_math,
//     -ftz=false), so results are bit-identical to the host loop c[i] = a[i] + b[i].
#include "common.cuh"
#include "gen.cuh"

namespace gp {

constexpr int VADD_THREADS = 256;

// ------------------------------------------------------------------ vector load / store

template <int VF>
struct Vec;  // VF floats per access
template <>
struct Vec<4> {
    float v[4];
};
template <>
struct Vec<8> {
    float v[8];
};

template <int VF, bool EVICT_FIRST>
__device__ __forceinline__ Vec<VF> vload(const float *p) {
    Vec<VF> r;
    if constexpr (VF == 4) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3])
                     : "l"(p));
    } else if constexpr (EVICT_FIRST) {
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]),
                       "=f"(r.v[6]), "=f"(r.v[7])
                     : "l"(p));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]),
                       "=f"(r.v[6]), "=f"(r.v[7])
                     : "l"(p));
    }
    return r;
}

template <int VF, bool EVICT_FIRST>
__device__ __forceinline__ void vstore(float *p, const Vec<VF> &r) {
    if constexpr (VF == 4) {
        asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]),
                     "f"(r.v[2]), "f"(r.v[3])
                     : "memory");
    } else if constexpr (EVICT_FIRST) {
        asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
                     "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]),
                     "f"(r.v[7])
                     : "memory");
    } else {
        asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.v[0]),
                     "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
                     : "memory");
    }
}

template <int VF>
__device__ __forceinline__ Vec<VF> vadd(const Vec<VF> &x, const Vec<VF> &y) {
    Vec<VF> r;
#pragma unroll
    for (int i = 0; i < VF; ++i) r.v[i] = __fadd_rn(x.v[i], y.v[i]);
    return r;
}

// One tile = VADD_THREADS * VF * UNROLL consecutive elements.  Thread t touches element
// tile*TILE + u*(THREADS*VF) + t*VF: each warp-wide access is one contiguous 512 B / 1 KiB run.
template <int VF, int UNROLL, bool EF>
__device__ __forceinline__ void vadd_one_tile(const float *__restrict__ a, const float *__restrict__ b,
                                              float *__restrict__ c, size_t tile) {
    constexpr size_t TILE = static_cast<size_t>(VADD_THREADS) * VF * UNROLL;
    constexpr size_t STEP = static_cast<size_t>(VADD_THREADS) * VF;
    const size_t base = tile * TILE + static_cast<size_t>(threadIdx.x) * VF;
    Vec<VF> va[UNROLL], vb[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) va[u] = vload<VF, EF>(a + base + u * STEP);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) vb[u] = vload<VF, EF>(b + base + u * STEP);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) vstore<VF, EF>(c + base + u * STEP, vadd<VF>(va[u], vb[u]));
}

// Non-persistent: one CTA per tile; grid = n_tiles.
template <int VF, int UNROLL, bool EF>
__global__ void __launch_bounds__(VADD_THREADS)
vadd_tile_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c) {
    vadd_one_tile<VF, UNROLL, EF>(a, b, c, blockIdx.x);
}

// Persistent: grid = k * SM count, CTAs walk tiles with a grid stride.
template <int VF, int UNROLL, bool EF>
__global__ void __launch_bounds__(VADD_THREADS)
vadd_persistent_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c,
                       uint64_t n_tiles) {
    for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) vadd_one_tile<VF, UNROLL, EF>(a, b, c, t);
}

// Any alignment, any length: 32-bit accesses with a bounds check (ragged tails, odd pointers).
__global__ void __launch_bounds__(VADD_THREADS)
vadd_scalar_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, uint64_t n) {
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        c[i] = __fadd_rn(__ldg(a + i), __ldg(b + i));
}

// Launch-latency-bound sizes (the sample's 50,000 elements): ONE launch, 128-bit accesses with a per-thread
// bounds check, the last < 4 elements by scalar accesses in the same kernel. Needs 16-byte aligned pointers.
__global__ void __launch_bounds__(VADD_THREADS)
vadd_small_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, uint32_t n) {
    const uint32_t i = (blockIdx.x * VADD_THREADS + threadIdx.x) * 4u;
    if (i + 3u < n) {
        const Vec<4> x = vload<4, false>(a + i), y = vload<4, false>(b + i);
        vstore<4, false>(c + i, vadd<4>(x, y));
    } else {
        for (uint32_t j = i; j < n; ++j) c[j] = __fadd_rn(__ldg(a + j), __ldg(b + j));
    }
}

// ------------------------------------------------------------------ TMA (cp.async.bulk) ring

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra W_%=;\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}

// Persistent CTAs stream TILE_F-float tiles of A and B into a STAGES-deep shared-memory ring with
// bulk async copies (mbarrier transaction-count completion), all threads add out of shared memory
// into the stage's C buffer, and one thread sends it back with a bulk store.
//   per iteration k (stage s = k % STAGES):
//     wait full[s]; LDS a,b; add; STS c[s]; fence.proxy.async; (thread 0: bulk wait_group.read STAGES-2);
//     __syncthreads(); thread 0: bulk store c[s], commit, refill a[s], b[s] with tile k+STAGES.
//   The wait_group.read before the barrier guarantees store k+1-STAGES has drained c[(k+1)%STAGES]
//   before anyone overwrites it in iteration k+1.
template <int TILE_F, int STAGES>
__global__ void __launch_bounds__(VADD_THREADS)
vadd_tma_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ c, uint64_t n_tiles) {
    static_assert(STAGES >= 3, "ring needs >= 3 stages (see wait_group.read)");
    constexpr uint32_t TILE_B = TILE_F * 4u;
    extern __shared__ __align__(128) unsigned char smem[];
    float *sa = reinterpret_cast<float *>(smem);
    float *sb = reinterpret_cast<float *>(smem + static_cast<size_t>(STAGES) * TILE_B);
    float *sc = reinterpret_cast<float *>(smem + 2u * static_cast<size_t>(STAGES) * TILE_B);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + 3u * static_cast<size_t>(STAGES) * TILE_B);

    const uint64_t first = blockIdx.x;
    const uint64_t stride = gridDim.x;
    const uint64_t my_tiles = first < n_tiles ? (n_tiles - first + stride - 1) / stride : 0;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto refill = [&](uint64_t k) {  // thread 0 only
        const int s = static_cast<int>(k % STAGES);
        const size_t off = (first + k * stride) * static_cast<size_t>(TILE_F);
        mbar_expect_tx(&full[s], 2u * TILE_B);
        bulk_g2s(sa + static_cast<size_t>(s) * TILE_F, a + off, TILE_B, &full[s]);
        bulk_g2s(sb + static_cast<size_t>(s) * TILE_F, b + off, TILE_B, &full[s]);
    };

    if (threadIdx.x == 0) {
        for (uint64_t k = 0; k < STAGES && k < my_tiles; ++k) refill(k);
    }

    constexpr int VPT = TILE_F / (VADD_THREADS * 4);  // float4 per thread per tile
    for (uint64_t k = 0; k < my_tiles; ++k) {
        const int s = static_cast<int>(k % STAGES);
        const uint32_t parity = static_cast<uint32_t>((k / STAGES) & 1u);
        mbar_wait(&full[s], parity);

        const float4 *pa = reinterpret_cast<const float4 *>(sa + static_cast<size_t>(s) * TILE_F);
        const float4 *pb = reinterpret_cast<const float4 *>(sb + static_cast<size_t>(s) * TILE_F);
        float4 *pc = reinterpret_cast<float4 *>(sc + static_cast<size_t>(s) * TILE_F);
#pragma unroll
        for (int v = 0; v < VPT; ++v) {
            const int idx = v * VADD_THREADS + threadIdx.x;
            const float4 x = pa[idx];
            const float4 y = pb[idx];
            float4 r;
            r.x = __fadd_rn(x.x, y.x);
            r.y = __fadd_rn(x.y, y.y);
            r.z = __fadd_rn(x.z, y.z);
            r.w = __fadd_rn(x.w, y.w);
            pc[idx] = r;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(STAGES - 2) : "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            const size_t off = (first + k * stride) * static_cast<size_t>(TILE_F);
            bulk_s2g(c + off, sc + static_cast<size_t>(s) * TILE_F, TILE_B);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (k + STAGES < my_tiles) refill(k + STAGES);
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ------------------------------------------------------------------ fill / digest / verify

// Generates elements [first, first+n) of A and B (either pointer may be null). Each thread owns
// four consecutive elements so aligned buffers get 128-bit stores.
__global__ void __launch_bounds__(256)
fill_inputs_kernel(float *__restrict__ a, float *__restrict__ b, uint64_t first, uint64_t n, uint32_t base_a,
                   uint32_t base_b, int kind, int vec_ok) {
    const uint64_t groups = (n + 3) / 4;
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t g = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; g < groups; g += stride) {
        const uint64_t j = g * 4;
        uint32_t ua[4], ub[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            ua[e] = gen_bits(base_a, first + j + e, kind, false);
            ub[e] = gen_bits(base_b, first + j + e, kind, true);
        }
        if (vec_ok && j + 3 < n) {
            if (a) stg128_stream(a + j, make_uint4(ua[0], ua[1], ua[2], ua[3]));
            if (b) stg128_stream(b + j, make_uint4(ub[0], ub[1], ub[2], ub[3]));
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (j + e < n) {
                    if (a) a[j + e] = __uint_as_float(ua[e]);
                    if (b) b[j + e] = __uint_as_float(ub[e]);
                }
            }
        }
    }
}

// w[i] = hash_at(base, i): source words of the P2P probe, per-rank inputs of the allreduce probe.
__global__ void __launch_bounds__(256) fill_words_kernel(uint32_t *__restrict__ w, uint64_t n, uint32_t base) {
    const uint64_t groups = (n + 3) / 4;
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t g = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; g < groups; g += stride) {
        const uint64_t j = g * 4;
        if (j + 3 < n) {
            stg128_stream(w + j, make_uint4(hash_at(base, j), hash_at(base, j + 1), hash_at(base, j + 2), hash_at(base, j + 3)));
        } else {
            for (uint64_t e = j; e < n; ++e) w[e] = hash_at(base, e);
        }
    }
}

// Digest scratch layout (device): one allocation carved as below.
struct DigestScratch {
    uint64_t out_sum;
    uint32_t out_xor;
    unsigned int ticket;  // must be 0 at launch; the last CTA resets it
    // followed by uint64_t partial_sum[max_blocks]; uint32_t partial_xor[max_blocks];
};
constexpr int DIGEST_MAX_BLOCKS = 148 * 8;
constexpr size_t DIGEST_SCRATCH_BYTES =
    sizeof(DigestScratch) + DIGEST_MAX_BLOCKS * (sizeof(uint64_t) + sizeof(uint32_t));

// sum64 / xor32 over n 32-bit words: 128-bit loads, warp-shuffle reduction, one partial per CTA,
// the last CTA to finish (ticket) folds the partials.
__global__ void __launch_bounds__(256)
digest_kernel(const uint32_t *__restrict__ w, uint64_t n, int vec_ok, DigestScratch *scr) {
    uint64_t *partial_sum = reinterpret_cast<uint64_t *>(scr + 1);
    uint32_t *partial_xor = reinterpret_cast<uint32_t *>(partial_sum + DIGEST_MAX_BLOCKS);
    uint64_t s = 0;
    uint32_t x = 0;
    const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    uint64_t done = 0;
    if (vec_ok) {
        const uint64_t n4 = n / 4;
        for (uint64_t i = tid; i < n4; i += stride) {
            const uint4 v = ldg128_stream(w + i * 4);
            s += static_cast<uint64_t>(v.x) + v.y + static_cast<uint64_t>(v.z) + v.w;
            x ^= v.x ^ v.y ^ v.z ^ v.w;
        }
        done = n4 * 4;
    }
    for (uint64_t i = done + tid; i < n; i += stride) {
        const uint32_t v = __ldg(w + i);
        s += v;
        x ^= v;
    }
    __shared__ uint64_t sh_s[8];
    __shared__ uint32_t sh_x[8];
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
        for (int i = 0; i < 8; ++i) {
            bs += sh_s[i];
            bx ^= sh_x[i];
        }
        partial_sum[blockIdx.x] = bs;
        partial_xor[blockIdx.x] = bx;
        __threadfence();
        const unsigned int t = atomicAdd(&scr->ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        uint64_t ts = 0;
        uint32_t tx = 0;
        for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
            ts += *(volatile uint64_t *)(partial_sum + i);
            tx ^= *(volatile uint32_t *)(partial_xor + i);
        }
        ts = warp_sum_u64(ts);
        tx = warp_xor_u32(tx);
        if (lane == 0) {
            sh_s[warp] = ts;
            sh_x[warp] = tx;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint64_t bs = 0;
            uint32_t bx = 0;
            for (int i = 0; i < 8; ++i) {
                bs += sh_s[i];
                bx ^= sh_x[i];
            }
            scr->out_sum = bs;
            scr->out_xor = bx;
            scr->ticket = 0;
        }
    }
}

// Counts elements of c that differ from regenerated a + b (same adder: a sanity check).
__global__ void __launch_bounds__(256)
verify_kernel(const float *__restrict__ c, uint64_t first, uint64_t n, uint32_t base_a, uint32_t base_b, int kind,
              unsigned long long *mismatches) {
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    unsigned long long bad = 0;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float fa = __uint_as_float(gen_bits(base_a, first + i, kind, false));
        const float fb = __uint_as_float(gen_bits(base_b, first + i, kind, true));
        bad += (__float_as_uint(__fadd_rn(fa, fb)) != __float_as_uint(__ldg(c + i)));
    }
    bad = warp_sum_u64(bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatches, bad);
}

// ------------------------------------------------------------------ host: dispatch

// Measured best on B200 at every HBM-resident size (profiles/r01_vadd_ncu_full.md, sections 3-6): the smallest tile.
constexpr int DEFAULT_VARIANT = GPUPROBE_VADD_LDG128_U1;
constexpr uint64_t SMALL_N_LIMIT = 1ull << 22;  // below this a launch costs more than the data movement

struct VariantInfo {
    int vec_floats;  // alignment requirement in floats (0: none)
    size_t tile;     // elements per tile (0: not tiled)
};

static VariantInfo variant_info(int v) {
    switch (v) {
    case GPUPROBE_VADD_LDG128_U1: return {4, 256u * 4 * 1};
    case GPUPROBE_VADD_LDG256_U1: return {8, 256u * 8 * 1};
    case GPUPROBE_VADD_LDG128_U2: return {4, 256u * 4 * 2};
    case GPUPROBE_VADD_LDG128_U4: return {4, 256u * 4 * 4};
    case GPUPROBE_VADD_LDG128_U8: return {4, 256u * 4 * 8};
    case GPUPROBE_VADD_LDG256_U2: return {8, 256u * 8 * 2};
    case GPUPROBE_VADD_LDG256_U4: return {8, 256u * 8 * 4};
    case GPUPROBE_VADD_LDG256_U2_EF: return {8, 256u * 8 * 2};
    case GPUPROBE_VADD_TMA: return {4, 4096};
    case GPUPROBE_VADD_TMA_8K: return {4, 2048};
    case GPUPROBE_VADD_PERSIST128: return {4, 256u * 4 * 4};
    default: return {0, 0};
    }
}

constexpr int TMA16_STAGES = 4, TMA16_TILE = 4096;
constexpr int TMA8_STAGES = 4, TMA8_TILE = 2048;
constexpr size_t tma_smem_bytes(int tile_f, int stages) { return 3u * stages * tile_f * 4u + stages * 8u; }

// Opt-in shared memory of the TMA variants: once per device, whichever thread gets there first (the device is current).
static int set_func_attrs(DeviceState &st, char *err, size_t errlen) {
    std::call_once(st.func_attrs_once, [&st]() {
        cudaError_t e = cudaFuncSetAttribute(vadd_tma_kernel<TMA16_TILE, TMA16_STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(tma_smem_bytes(TMA16_TILE, TMA16_STAGES)));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(vadd_tma_kernel<TMA8_TILE, TMA8_STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(tma_smem_bytes(TMA8_TILE, TMA8_STAGES)));
        st.func_attrs_err = e;
    });
    if (st.func_attrs_err != cudaSuccess) {
        set_err(err, errlen, "cudaFuncSetAttribute failed: %s", cudaGetErrorString(st.func_attrs_err));
        (void)cudaGetLastError();
        return GPUPROBE_ECUDA;
    }
    return GPUPROBE_OK;
}

static int scalar_grid(const DeviceState &st, uint64_t n) {
    const uint64_t want = (n + VADD_THREADS - 1) / VADD_THREADS;
    const uint64_t cap = static_cast<uint64_t>(st.sm_count) * 8;
    return static_cast<int>(std::max<uint64_t>(1, std::min(want, cap)));
}

// Enqueues C = A + B on `stream`. Returns the number of kernel launches (>= 0) or an error (< 0).
int launch_vadd(DeviceState &st, const float *a, const float *b, float *c, uint64_t n, int variant,
                cudaStream_t stream, char *err, size_t errlen) {
    if (n == 0) return 0;
    if (variant < 0 || variant >= GPUPROBE_VADD_NUM_VARIANTS) {
        set_err(err, errlen, "unknown vectorAdd variant %d", variant);
        return GPUPROBE_EINVAL;
    }
    const bool aligned16 = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                             reinterpret_cast<uintptr_t>(c)) & 15u) == 0;
    if (variant == GPUPROBE_VADD_SMALL || (variant == GPUPROBE_VADD_AUTO && n < SMALL_N_LIMIT && n % 1024u != 0)) {
        // a ragged launch-bound size would otherwise cost two launches (tiles + tail)
        if (aligned16 && n < (1ull << 31)) {
            const unsigned grid = static_cast<unsigned>((n + VADD_THREADS * 4 - 1) / (VADD_THREADS * 4));
            vadd_small_kernel<<<grid, VADD_THREADS, 0, stream>>>(a, b, c, static_cast<uint32_t>(n));
            GP_CUDA(cudaGetLastError());
            return 1;
        }
        if (variant == GPUPROBE_VADD_SMALL) variant = GPUPROBE_VADD_SCALAR;
    }
    if (variant == GPUPROBE_VADD_AUTO) variant = DEFAULT_VARIANT;
    VariantInfo vi = variant_info(variant);
    const uintptr_t align_mask = vi.vec_floats ? static_cast<uintptr_t>(vi.vec_floats) * 4u - 1u : 0u;
    const bool aligned = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                           reinterpret_cast<uintptr_t>(c)) & align_mask) == 0;
    uint64_t n_tiles = (vi.tile && aligned) ? n / vi.tile : 0;
    int launches = 0;

    if (n_tiles > 0) {
        if (n_tiles > 0x7fffffffULL && variant != GPUPROBE_VADD_TMA && variant != GPUPROBE_VADD_TMA_8K &&
            variant != GPUPROBE_VADD_PERSIST128) {
            set_err(err, errlen, "n too large for a one-CTA-per-tile grid (%llu tiles)", (unsigned long long)n_tiles);
            return GPUPROBE_EINVAL;
        }
        const dim3 grid(static_cast<unsigned>(n_tiles)), block(VADD_THREADS);
        switch (variant) {
        case GPUPROBE_VADD_LDG128_U1: vadd_tile_kernel<4, 1, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG256_U1: vadd_tile_kernel<8, 1, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG128_U2: vadd_tile_kernel<4, 2, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG128_U4: vadd_tile_kernel<4, 4, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG128_U8: vadd_tile_kernel<4, 8, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG256_U2: vadd_tile_kernel<8, 2, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG256_U4: vadd_tile_kernel<8, 4, false><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_LDG256_U2_EF: vadd_tile_kernel<8, 2, true><<<grid, block, 0, stream>>>(a, b, c); break;
        case GPUPROBE_VADD_PERSIST128: {
            const unsigned g = static_cast<unsigned>(std::min<uint64_t>(n_tiles, static_cast<uint64_t>(st.sm_count) * 8));
            vadd_persistent_kernel<4, 4, false><<<g, block, 0, stream>>>(a, b, c, n_tiles);
            break;
        }
        case GPUPROBE_VADD_TMA: {
            int rc = set_func_attrs(st, err, errlen);
            if (rc) return rc;
            const unsigned g = static_cast<unsigned>(std::min<uint64_t>(n_tiles, static_cast<uint64_t>(st.sm_count)));
            vadd_tma_kernel<TMA16_TILE, TMA16_STAGES>
                <<<g, block, tma_smem_bytes(TMA16_TILE, TMA16_STAGES), stream>>>(a, b, c, n_tiles);
            break;
        }
        case GPUPROBE_VADD_TMA_8K: {
            int rc = set_func_attrs(st, err, errlen);
            if (rc) return rc;
            const unsigned g = static_cast<unsigned>(std::min<uint64_t>(n_tiles, static_cast<uint64_t>(st.sm_count) * 2));
            vadd_tma_kernel<TMA8_TILE, TMA8_STAGES>
                <<<g, block, tma_smem_bytes(TMA8_TILE, TMA8_STAGES), stream>>>(a, b, c, n_tiles);
            break;
        }
        default: break;
        }
        GP_CUDA(cudaGetLastError());
        ++launches;
    }
    const uint64_t done = n_tiles * vi.tile;
    if (done < n) {
        const uint64_t rem = n - done;
        vadd_scalar_kernel<<<scalar_grid(st, rem), VADD_THREADS, 0, stream>>>(a + done, b + done, c + done, rem);
        GP_CUDA(cudaGetLastError());
        ++launches;
    }
    return launches;
}

static int launch_fill(DeviceState &st, float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int kind,
                       cudaStream_t stream, char *err, size_t errlen) {
    if (n == 0 || (!a && !b)) return GPUPROBE_OK;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15u) == 0;
    const uint64_t groups = (n + 3) / 4;
    const uint64_t want = (groups + 255) / 256;
    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(st.sm_count) * 16)));
    fill_inputs_kernel<<<grid, 256, 0, stream>>>(a, b, first, n, hash_base(seed, STREAM_A), hash_base(seed, STREAM_B),
                                                 kind, vec_ok);
    GP_CUDA(cudaGetLastError());
    return GPUPROBE_OK;
}

int launch_fill_words(DeviceState &st, uint32_t *d_w, uint64_t n, uint32_t hash_base_value, cudaStream_t stream,
                      char *err, size_t errlen) {
    if (n == 0) return GPUPROBE_OK;
    if (reinterpret_cast<uintptr_t>(d_w) & 15u) {
        set_err(err, errlen, "fill_words: buffer must be 16-byte aligned");
        return GPUPROBE_EINVAL;
    }
    const uint64_t want = ((n + 3) / 4 + 255) / 256;
    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(st.sm_count) * 16)));
    fill_words_kernel<<<grid, 256, 0, stream>>>(d_w, n, hash_base_value);
    GP_CUDA(cudaGetLastError());
    return GPUPROBE_OK;
}

static int digest_grid(const DeviceState &st, uint64_t n) {
    const uint64_t want = (n / 4 + 255) / 256;
    const uint64_t cap = std::min<uint64_t>(static_cast<uint64_t>(st.sm_count) * 8, DIGEST_MAX_BLOCKS);
    return static_cast<int>(std::max<uint64_t>(1, std::min(want, cap)));
}

// The per-device scratch of the digest and verify kernels: [DigestScratch + partials][mismatch counter]. Allocated at the
// first use on the device and kept (gpuprobe_shutdown frees it), so that neither the 50,000-element drop-in probe nor a
// digest inside a timed pass pays a cudaMalloc/cudaFree pair (both take device-wide locks). Caller holds st.scratch_mu.
constexpr size_t VERIFY_COUNTER_OFFSET = (DIGEST_SCRATCH_BYTES + 15) & ~static_cast<size_t>(15);
constexpr size_t DEVICE_SCRATCH_BYTES = VERIFY_COUNTER_OFFSET + 16;

static int ensure_scratch(DeviceState &st, char *err, size_t errlen) {
    if (st.scratch) return GPUPROBE_OK;
    GP_CUDA(cudaMalloc(&st.scratch, DEVICE_SCRATCH_BYTES));
    return GPUPROBE_OK;
}

// Digest of n words at d_words; uses the device's cached scratch and synchronises the stream.
int run_digest(DeviceState &st, const void *d_words, uint64_t n, uint64_t *sum64, uint32_t *xor32,
                      cudaStream_t stream, char *err, size_t errlen) {
    if (n == 0) {
        *sum64 = 0;
        *xor32 = 0;
        return GPUPROBE_OK;
    }
    std::lock_guard<std::mutex> lock(st.scratch_mu);
    int rc = ensure_scratch(st, err, errlen);
    if (rc) return rc;
    DigestScratch *scr = static_cast<DigestScratch *>(st.scratch);
    DigestScratch host{};
    // result words and ticket start at zero (a kernel that died mid-way may have left a stale ticket)
    cudaError_t e = cudaMemsetAsync(scr, 0, sizeof(DigestScratch), stream);
    if (e == cudaSuccess) {
        const int vec_ok = (reinterpret_cast<uintptr_t>(d_words) & 15u) == 0;
        digest_kernel<<<digest_grid(st, n), 256, 0, stream>>>(static_cast<const uint32_t *>(d_words), n, vec_ok, scr);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(&host, scr, sizeof(DigestScratch), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        set_err(err, errlen, "digest failed: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return GPUPROBE_ECUDA;
    }
    *sum64 = host.out_sum;
    *xor32 = host.out_xor;
    return GPUPROBE_OK;
}

static int run_verify(DeviceState &st, const float *d_c, uint64_t first, uint64_t n, uint32_t seed, int kind,
                      uint64_t *mismatches, cudaStream_t stream, char *err, size_t errlen) {
    *mismatches = 0;
    if (n == 0) return GPUPROBE_OK;
    std::lock_guard<std::mutex> lock(st.scratch_mu);
    int rc = ensure_scratch(st, err, errlen);
    if (rc) return rc;
    unsigned long long *d_bad = reinterpret_cast<unsigned long long *>(static_cast<unsigned char *>(st.scratch) + VERIFY_COUNTER_OFFSET);
    unsigned long long h_bad = 0;
    cudaError_t e = cudaMemsetAsync(d_bad, 0, sizeof(unsigned long long), stream);
    if (e == cudaSuccess) {
        verify_kernel<<<scalar_grid(st, n), 256, 0, stream>>>(d_c, first, n, hash_base(seed, STREAM_A),
                                                              hash_base(seed, STREAM_B), kind, d_bad);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h_bad, d_bad, sizeof(h_bad), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        set_err(err, errlen, "verify failed: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return GPUPROBE_ECUDA;
    }
    *mismatches = h_bad;
    return GPUPROBE_OK;
}

static bool valid_kind(int k) { return k >= GPUPROBE_GEN_UNIT && k <= GPUPROBE_GEN_BITS; }

}  // namespace gp

// ====================================================================== C ABI

using namespace gp;

extern "C" int gpuprobe_vectoradd_device(int dev, const float *d_a, const float *d_b, float *d_c, uint64_t n,
                                         int variant, void *stream, char *err, size_t errlen) {
    DeviceState *st = nullptr;
    int rc = use_device(dev, &st, err, errlen);
    if (rc) return rc;
    if (n && (!d_a || !d_b || !d_c)) {
        set_err(err, errlen, "null device pointer");
        return GPUPROBE_EINVAL;
    }
    rc = launch_vadd(*st, d_a, d_b, d_c, n, variant, static_cast<cudaStream_t>(stream), err, errlen);
    return rc < 0 ? rc : GPUPROBE_OK;
}

extern "C" int gpuprobe_fill_inputs_device(int dev, float *d_a, float *d_b, uint64_t first, uint64_t n, uint32_t seed,
                                           int gen_kind, void *stream, char *err, size_t errlen) {
    DeviceState *st = nullptr;
    int rc = use_device(dev, &st, err, errlen);
    if (rc) return rc;
    if (!valid_kind(gen_kind)) {
        set_err(err, errlen, "unknown gen_kind %d", gen_kind);
        return GPUPROBE_EINVAL;
    }
    return launch_fill(*st, d_a, d_b, first, n, seed, gen_kind, static_cast<cudaStream_t>(stream), err, errlen);
}

extern "C" int gpuprobe_digest_device(int dev, const void *d_words, uint64_t n, uint64_t *sum64, uint32_t *xor32,
                                      void *stream, char *err, size_t errlen) {
    DeviceState *st = nullptr;
    int rc = use_device(dev, &st, err, errlen);
    if (rc) return rc;
    if (!sum64 || !xor32 || (n && !d_words)) {
        set_err(err, errlen, "null pointer");
        return GPUPROBE_EINVAL;
    }
    return run_digest(*st, d_words, n, sum64, xor32, static_cast<cudaStream_t>(stream), err, errlen);
}

extern "C" int gpuprobe_verify_device(int dev, const float *d_c, uint64_t first, uint64_t n, uint32_t seed,
                                      int gen_kind, uint64_t *mismatches, void *stream, char *err, size_t errlen) {
    DeviceState *st = nullptr;
    int rc = use_device(dev, &st, err, errlen);
    if (rc) return rc;
    if (!mismatches || (n && !d_c) || !valid_kind(gen_kind)) {
        set_err(err, errlen, "bad argument");
        return GPUPROBE_EINVAL;
    }
    return run_verify(*st, d_c, first, n, seed, gen_kind, mismatches, static_cast<cudaStream_t>(stream), err, errlen);
}

extern "C" int gpuprobe_vectoradd(const gpuprobe_vadd_cfg *cfg, gpuprobe_vadd_result *out, char *err, size_t errlen) {
    if (!cfg || !out) {
        set_err(err, errlen, "null cfg/out");
        return GPUPROBE_EINVAL;
    }
    std::memset(out, 0, sizeof(*out));
    if (cfg->iters < 1 || cfg->warmup < 0 || !valid_kind(cfg->gen_kind) || cfg->variant < 0 ||
        cfg->variant >= GPUPROBE_VADD_NUM_VARIANTS) {
        set_err(err, errlen, "bad cfg: iters=%d warmup=%d gen_kind=%d variant=%d", cfg->iters, cfg->warmup,
                cfg->gen_kind, cfg->variant);
        return GPUPROBE_EINVAL;
    }
    DeviceState *stp = nullptr;
    int rc = use_device(cfg->dev, &stp, err, errlen);
    if (rc) return rc;
    DeviceState &st = *stp;
    std::lock_guard<std::mutex> lock(st.mu);

    const uint64_t n = cfg->n_elems;
    out->n = n;
    out->algo_bytes = 12ull * n;
    out->variant = static_cast<uint32_t>(cfg->variant != GPUPROBE_VADD_AUTO           ? cfg->variant
                                         : (cfg->n_elems < SMALL_N_LIMIT && cfg->n_elems % 1024u) ? GPUPROBE_VADD_SMALL
                                                                                                  : DEFAULT_VARIANT);
    out->l2_resident = (12ull * n <= st.l2_bytes) ? 1 : 0;
    if (n == 0) return GPUPROBE_OK;  // empty input: nothing to add, trivially correct

    float *d_a = nullptr, *d_b = nullptr, *d_c = nullptr;
    std::vector<cudaEvent_t> ev;
    auto cleanup = [&]() {
        for (cudaEvent_t e : ev) cudaEventDestroy(e);
        cudaFree(d_a);
        cudaFree(d_b);
        cudaFree(d_c);
    };
    auto fail = [&](int code) {
        cleanup();
        return code;
    };
#define GP_TRY(call)                                                                                   \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            set_err(err, errlen, "%s failed: %s", #call, cudaGetErrorString(e_));                      \
            (void)cudaGetLastError();                                                                  \
            return fail(e_ == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA);           \
        }                                                                                              \
    } while (0)

    GP_TRY(cudaMalloc(&d_a, n * sizeof(float)));
    GP_TRY(cudaMalloc(&d_b, n * sizeof(float)));
    GP_TRY(cudaMalloc(&d_c, n * sizeof(float)));
    rc = launch_fill(st, d_a, d_b, 0, n, cfg->seed, cfg->gen_kind, st.stream, err, errlen);
    if (rc) return fail(rc);
    // poison C so that a kernel that writes nothing cannot pass on stale data
    GP_TRY(cudaMemsetAsync(d_c, 0xff, n * sizeof(float), st.stream));

    int launches = 0;
    for (int i = 0; i < cfg->warmup; ++i) {
        launches = launch_vadd(st, d_a, d_b, d_c, n, cfg->variant, st.stream, err, errlen);
        if (launches < 0) return fail(launches);
    }
    ev.resize(2 * static_cast<size_t>(cfg->iters));
    for (auto &e : ev) e = nullptr;
    for (auto &e : ev) GP_TRY(cudaEventCreate(&e));
    for (int i = 0; i < cfg->iters; ++i) {
        GP_TRY(cudaEventRecord(ev[2 * i], st.stream));
        launches = launch_vadd(st, d_a, d_b, d_c, n, cfg->variant, st.stream, err, errlen);
        if (launches < 0) return fail(launches);
        GP_TRY(cudaEventRecord(ev[2 * i + 1], st.stream));
    }
    GP_TRY(cudaStreamSynchronize(st.stream));
    std::vector<float> ms(cfg->iters);
    for (int i = 0; i < cfg->iters; ++i) GP_TRY(cudaEventElapsedTime(&ms[i], ev[2 * i], ev[2 * i + 1]));
    // ms_mean: the same number of launches back to back inside ONE event pair. For launch-bound sizes (the
    // sample's 50,000 elements) this is the per-launch cost without the event records in between.
    GP_TRY(cudaEventRecord(ev[0], st.stream));
    for (int i = 0; i < cfg->iters; ++i) {
        launches = launch_vadd(st, d_a, d_b, d_c, n, cfg->variant, st.stream, err, errlen);
        if (launches < 0) return fail(launches);
    }
    GP_TRY(cudaEventRecord(ev[1], st.stream));
    GP_TRY(cudaStreamSynchronize(st.stream));
    float batch_ms = 0;
    GP_TRY(cudaEventElapsedTime(&batch_ms, ev[0], ev[1]));
    out->launches = launches;
    out->ms_best = *std::min_element(ms.begin(), ms.end());
    out->ms_median = median_of(ms);
    out->ms_mean = batch_ms / cfg->iters;
    out->gbps_best = out->ms_best > 0 ? out->algo_bytes / (out->ms_best * 1e6) : 0;
    out->gbps_median = out->ms_median > 0 ? out->algo_bytes / (out->ms_median * 1e6) : 0;
    const double nominal = cfg->peak_nominal_gbps > 0 ? cfg->peak_nominal_gbps : 8000.0;
    out->frac_nominal = out->gbps_median / nominal;
    out->frac_measured = cfg->peak_measured_gbps > 0 ? out->gbps_median / cfg->peak_measured_gbps : 0;

    if (cfg->verify & 1) {
        rc = run_digest(st, d_c, n, &out->sum64, &out->xor32, st.stream, err, errlen);
        if (rc) return fail(rc);
    }
    if (cfg->verify & 2) {
        rc = run_verify(st, d_c, 0, n, cfg->seed, cfg->gen_kind, &out->mismatches, st.stream, err, errlen);
        if (rc) return fail(rc);
        if (out->mismatches) {
            set_err(err, errlen, "vectorAdd: %llu of %llu elements differ from a+b", (unsigned long long)out->mismatches,
                    (unsigned long long)n);
            return fail(GPUPROBE_FAIL_VERIFY);
        }
    }
    cleanup();
    return GPUPROBE_OK;
#undef GP_TRY
}

// ---------------------------------------------------------------- host buffers (the e2e path)

namespace gp {
constexpr uint64_t BUF_CHUNK_ELEMS = 8ull << 20;  // default: 32 MiB per array per chunk
constexpr int BUF_SLOTS = 3;                      // default number of chunks in flight

// Makes slot `sl` able to stage `chunk` elements (and to bounce them through pinned memory when `bounce`). The buffers
// are cached in DeviceState and only ever grow: the drop-in vectorAdd (50,000 elements, the reference's one real
// workload) and a validator that re-probes periodically pay the allocations once per process, not per call.
static cudaError_t ensure_slot(BufSlot &sl, uint64_t chunk, bool bounce) {
    cudaError_t e = cudaSuccess;
    if (!sl.done && (e = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming)) != cudaSuccess) return e;
    if (sl.dev_cap < chunk) {
        cudaFree(sl.d_a);
        cudaFree(sl.d_b);
        cudaFree(sl.d_c);
        sl.d_a = sl.d_b = sl.d_c = nullptr;
        sl.dev_cap = 0;
        if ((e = cudaMalloc(&sl.d_a, chunk * sizeof(float))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&sl.d_b, chunk * sizeof(float))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&sl.d_c, chunk * sizeof(float))) != cudaSuccess) return e;
        sl.dev_cap = chunk;
    }
    if (bounce && sl.host_cap < chunk) {
        if (sl.h_a) cudaFreeHost(sl.h_a);
        if (sl.h_b) cudaFreeHost(sl.h_b);
        if (sl.h_c) cudaFreeHost(sl.h_c);
        sl.h_a = sl.h_b = sl.h_c = nullptr;
        sl.host_cap = 0;
        if ((e = cudaHostAlloc(&sl.h_a, chunk * sizeof(float), cudaHostAllocDefault)) != cudaSuccess) return e;
        if ((e = cudaHostAlloc(&sl.h_b, chunk * sizeof(float), cudaHostAllocDefault)) != cudaSuccess) return e;
        if ((e = cudaHostAlloc(&sl.h_c, chunk * sizeof(float), cudaHostAllocDefault)) != cudaSuccess) return e;
        sl.host_cap = chunk;
    }
    sl.pending_off = sl.pending_n = 0;
    return cudaSuccess;
}

static bool is_pinned(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// The address the device can use for a pinned host pointer (UVA), or null if the memory is not mapped.
static const void *device_alias(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return nullptr;
    }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}
}  // namespace gp

extern "C" int gpuprobe_vectoradd_buffers(int dev, const float *a, const float *b, float *c, uint64_t n, char *err,
                                          size_t errlen) {
    DeviceState *stp = nullptr;
    int rc = use_device(dev, &stp, err, errlen);
    if (rc) return rc;
    if (n == 0) return GPUPROBE_OK;
    if (!a || !b || !c) {
        set_err(err, errlen, "null host pointer");
        return GPUPROBE_EINVAL;
    }
    DeviceState &st = *stp;
    std::lock_guard<std::mutex> lock(st.mu);

    const bool direct = is_pinned(a) && is_pinned(b) && is_pinned(c);
    // Pinned, device-mapped buffers: no staging at all. The add kernel loads A and B straight from host memory over
    // PCIe and stores C back over PCIe, so both directions of the link run for the whole call and nothing is
    // allocated on the device. Measured at the link limit for every kernel variant (profiles/r01_e2e_chunks.md).
    // GPUPROBE_BUF_ZEROCOPY=0 forces the staged pipeline below (comparison runs).
    const char *zc = getenv("GPUPROBE_BUF_ZEROCOPY");
    if (direct && !(zc && *zc == '0')) {
        const float *da = static_cast<const float *>(device_alias(a)), *db = static_cast<const float *>(device_alias(b));
        float *dc = static_cast<float *>(const_cast<void *>(device_alias(c)));
        if (da && db && dc) {
            rc = launch_vadd(st, da, db, dc, n, GPUPROBE_VADD_AUTO, st.stream, err, errlen);
            if (rc < 0) return rc;
            GP_CUDA(cudaStreamSynchronize(st.stream));
            return GPUPROBE_OK;
        }
    }
    // tuning knobs for experiments; the defaults are the measured best (profiles/r01_e2e_chunks.md)
    uint64_t chunk_elems = BUF_CHUNK_ELEMS;
    int max_slots = BUF_SLOTS;
    if (const char *e = getenv("GPUPROBE_BUF_CHUNK_LOG2")) chunk_elems = 1ull << std::max(10, std::min(30, atoi(e)));
    if (const char *e = getenv("GPUPROBE_BUF_SLOTS")) max_slots = std::max(1, std::min(BUF_MAX_SLOTS, atoi(e)));
    const uint64_t chunk = std::min<uint64_t>(n, chunk_elems);
    const int slots = static_cast<int>(std::min<uint64_t>(static_cast<uint64_t>(max_slots), (n + chunk - 1) / chunk));
    BufSlot *slot = st.buf_slots;  // cached staging, guarded by st.mu (held)
    // After a failure the cache is dropped: whatever state the slots were in, the next call starts from nothing.
    auto cleanup = [&]() { free_buf_slots(st); };
#define GP_TRY(call)                                                                                   \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            set_err(err, errlen, "%s failed: %s", #call, cudaGetErrorString(e_));                      \
            (void)cudaGetLastError();                                                                  \
            cudaDeviceSynchronize();                                                                   \
            cleanup();                                                                                 \
            return e_ == cudaErrorMemoryAllocation ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;                 \
        }                                                                                              \
    } while (0)

    for (int s = 0; s < slots; ++s) GP_TRY(ensure_slot(slot[s], chunk, !direct));

    uint64_t off = 0;
    for (uint64_t k = 0; off < n; ++k, off += chunk) {
        const uint64_t m = std::min<uint64_t>(chunk, n - off);
        BufSlot &sl = slot[k % slots];
        cudaStream_t s = st.copy_streams[k % slots];
        const float *src_a = a + off, *src_b = b + off;
        float *dst_c = c + off;
        if (!direct) {
            // the slot's previous chunk must have landed in h_c before the bounce buffers are reused
            if (sl.pending_n) {
                GP_TRY(cudaEventSynchronize(sl.done));
                std::memcpy(c + sl.pending_off, sl.h_c, sl.pending_n * sizeof(float));
                sl.pending_n = 0;
            }
            std::memcpy(sl.h_a, src_a, m * sizeof(float));
            std::memcpy(sl.h_b, src_b, m * sizeof(float));
            src_a = sl.h_a;
            src_b = sl.h_b;
            dst_c = sl.h_c;
        }
        GP_TRY(cudaMemcpyAsync(sl.d_a, src_a, m * sizeof(float), cudaMemcpyHostToDevice, s));
        GP_TRY(cudaMemcpyAsync(sl.d_b, src_b, m * sizeof(float), cudaMemcpyHostToDevice, s));
        rc = launch_vadd(st, sl.d_a, sl.d_b, sl.d_c, m, GPUPROBE_VADD_AUTO, s, err, errlen);
        if (rc < 0) {
            cudaDeviceSynchronize();
            cleanup();
            return rc;
        }
        GP_TRY(cudaMemcpyAsync(dst_c, sl.d_c, m * sizeof(float), cudaMemcpyDeviceToHost, s));
        if (!direct) {
            GP_TRY(cudaEventRecord(sl.done, s));
            sl.pending_off = off;
            sl.pending_n = m;
        }
    }
    for (int s = 0; s < slots; ++s) GP_TRY(cudaStreamSynchronize(st.copy_streams[s]));
    if (!direct) {
        for (int s = 0; s < slots; ++s)
            if (slot[s].pending_n) std::memcpy(c + slot[s].pending_off, slot[s].h_c, slot[s].pending_n * sizeof(float));
    }
    // GPUPROBE_BUF_CACHE=0: give the staging back after every call (callers that probe once and then run something else)
    if (const char *e = getenv("GPUPROBE_BUF_CACHE"); e && *e == '0') cleanup();
    return GPUPROBE_OK;
#undef GP_TRY
}

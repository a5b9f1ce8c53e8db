// nvls.cu -- probe (c'): int32 allreduce as ONE kernel per GPU over NVLink multicast (NVLS).
//
// The NCCL probe (allreduce.cu) validates the library path. This one exercises the switch itself:
// every GPU's buffer is bound to one multicast object, and a single cooperative kernel per GPU does
//
//     fill (optional)  x_r[i] = (int32)hash(seed, AR(r), i)          local stores
//     barrier          multimem.red.release.sys on a flag word + ld.acquire.sys spin
//     reduce+broadcast for i in my 1/n slice:  v = multimem.ld_reduce.add.u32 [mc + i]   (NVSwitch pulls the word from
//                                               all n GPUs and adds in the switch)
//                                              multimem.st.u32 [mc + i], v                (NVSwitch writes v to all n GPUs)
//     barrier
//     verify (optional) compare every local word with mult * sum_r hash(seed, AR(r), i)
//
// so the transfer and the arithmetic are the same instructions; there is no separate compute step and no
// staging copy. Per GPU roughly count*4*(1 + 1/n) bytes leave and the same arrive, against 2(n-1)/n * count*4 each
// way for a ring, which is why busbw (= algbw * 2(n-1)/n) can exceed the 900 GB/s link figure.
//
// The same launch structure (one cooperative kernel per GPU, buffers mapped on every rank, flag barriers with a timeout)
// carries the other kernels of this file, selected by `algo`:
//     1  two-shot over peer mappings       pull-reduce my slice from every rank, push the sum to every rank
//     2  hybrid                            head of the slice through the switch, tail over peer mappings, disjoint CTAs
//     3  all-push two-shot                 scatter to the owner's staging area, add out of local HBM, broadcast: stores only
//     4  NVLS + copy engines               this kernel on the head of the buffer, ce_twoshot.cuh on the tail (measured: no gain)
// and gpuprobe_allreduce_i32_auto picks among them and the copy-engine two-shot by rank count and size.
//
// The reference has no counterpart.
//
// Driver-API calls (cuMem*, cuMulticast*) are resolved with cudaGetDriverEntryPoint, so the library still has
// no link-time dependency on libcuda. Cross-GPU spins have a timeout and raise an error flag instead of hanging.
#include <cooperative_groups.h>
#include <cuda.h>

#include "ce_twoshot.cuh"
#include "common.cuh"
#include "gen.cuh"

namespace cg = cooperative_groups;

namespace gp {

constexpr int NVLS_THREADS = 512;
constexpr uint64_t NVLS_FLAG_BYTES = 4096;  // the flag word lives in its own page-slice behind the payload
constexpr unsigned long long NVLS_SPIN_TIMEOUT_NS = 4000000000ull;

constexpr int NVLS_MAX_PEERS = 16;

enum : int { NVLS_FILL = 1, NVLS_REDUCE = 2, NVLS_VERIFY = 4 };

struct NvlsArgs {
    uint32_t *local;        // unicast mapping of this rank's buffer
    uint32_t *mc;           // multicast mapping of the same bytes on every rank
    uint32_t *flag_local;   // unicast address of the barrier word in this rank's memory
    uint32_t *flag_mc;      // multicast address of the barrier word
    unsigned long long *mismatches;
    int *error;             // set to 1 if a cross-GPU wait timed out
    uint64_t count;
    uint32_t seed;
    uint32_t mult;          // verify: expected = mult * sum_r x_r[i]
    uint32_t barrier_base;  // number of barriers completed before this launch
    int rank, nranks, phases;
    int algo;               // 0: NVLS multimem; 1: two-shot over peer mappings (pull-reduce my slice, push it to everyone);
                            // 2: hybrid (head of the slice by NVLS, tail by peer mappings, on disjoint CTAs);
                            // 3: all-push two-shot (scatter my copy of every slice into its owner's staging area with
                            //    peer stores, owner adds locally, broadcasts the sum with peer stores: no read crosses a link)
    int hybrid_permille;    // algo 2: share of the slice that goes through the switch, in 1/1000
    int hybrid_nvls_blocks; // algo 2: CTAs [0, this) run the NVLS part, the rest the peer part
    int barrier_peer;       // 1: no multicast on this box -- barrier through per-rank flag slots written with peer stores
    int st_v4;              // NVLS part: broadcast with 16-byte multimem.st (four sums per store) instead of 4-byte ones
    uint64_t stage_off;     // algo 3: word offset of the staging area (nranks-1 slices) behind the payload of every rank
    uint64_t reduce_count;  // words [0, reduce_count) are reduced by this kernel (= count, except in the NVLS + copy-engine hybrid)
    uint32_t *peer[NVLS_MAX_PEERS];  // algo 1: unicast mapping of every rank's buffer (peer[rank] == local)
};

__device__ __forceinline__ uint32_t mm_ld_reduce_add_u32(const uint32_t *mc) {
    uint32_t v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_u32(uint32_t *mc, uint32_t v) {
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
// Four 32-bit sums in one 16-byte multicast store. The .f32 type only names the container: a store does no
// arithmetic, the bit patterns travel unchanged (multimem.ld_reduce has no vector form for integer adds).
__device__ __forceinline__ void mm_st_v4(uint32_t *mc, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(__uint_as_float(x)),
                 "f"(__uint_as_float(y)), "f"(__uint_as_float(z)), "f"(__uint_as_float(w))
                 : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ uint4 ld_cg128(const uint32_t *p) {  // L2-coherent load (peer data changes between rounds)
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_128(uint32_t *p, const uint4 &v) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Two-shot allreduce of words [lo, hi) over peer mappings: load the slice from every rank (NR-1 loads cross NVLink),
// add, store the sum into every rank's buffer (NR-1 stores cross NVLink). NR is a compile-time rank count so that all
// peer loads of a vector are in flight together.
template <int NR>
__device__ __forceinline__ void p2p_reduce_broadcast(const NvlsArgs &a, uint64_t lo, uint64_t hi, uint64_t gtid, uint64_t gstride) {
    const uint64_t lo4 = lo / 4, hi4 = hi / 4;  // lo is a multiple of 32 words
    for (uint64_t v = lo4 + gtid; v < hi4; v += gstride) {
        uint4 x[NR];
#pragma unroll
        for (int r = 0; r < NR; ++r) x[r] = ld_cg128(a.peer[r] + v * 4);
        uint4 acc = x[0];
#pragma unroll
        for (int r = 1; r < NR; ++r) {
            acc.x += x[r].x;
            acc.y += x[r].y;
            acc.z += x[r].z;
            acc.w += x[r].w;
        }
#pragma unroll
        for (int r = 0; r < NR; ++r) st_128(a.peer[r] + v * 4, acc);
    }
    // < 4 trailing words of the whole buffer; max(): an empty slice clamped to `count` must not redo its neighbour's tail
    for (uint64_t i = max(hi4 * 4, lo) + gtid; i < hi; i += gstride) {
        uint32_t acc = 0;
        for (int r = 0; r < NR; ++r) {
            uint32_t w;
            asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(w) : "l"(a.peer[r] + i) : "memory");
            acc += w;
        }
        for (int r = 0; r < NR; ++r) a.peer[r][i] = acc;
    }
}

__device__ __forceinline__ uint32_t ld_cg32(const uint32_t *p) {
    uint32_t w;
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(w) : "l"(p) : "memory");
    return w;
}

// All-push two-shot, phase A: for every other rank q, copy MY words of q's slice into the staging slot q keeps for
// me, with peer stores. Loads are local HBM reads; only posted writes cross NVLink. `per` = slice length (multiple
// of 32 words); slice q = [q*per, min((q+1)*per, count)); rank r's slot at owner q is slot (r < q ? r : r-1).
template <int NR>
__device__ __forceinline__ void push_scatter(const NvlsArgs &a, uint64_t per, uint64_t gtid, uint64_t gstride) {
    const uint64_t per4 = per / 4, count4 = a.count / 4;
    for (uint64_t v = gtid; v < per4; v += gstride) {
        uint4 x[NR];
#pragma unroll
        for (int q = 0; q < NR; ++q) {
            const uint64_t w4 = static_cast<uint64_t>(q) * per4 + v;
            if (q != a.rank && w4 < count4) x[q] = ld_cg128(a.local + w4 * 4);
        }
#pragma unroll
        for (int q = 0; q < NR; ++q) {
            const uint64_t w4 = static_cast<uint64_t>(q) * per4 + v;
            const int slot = a.rank < q ? a.rank : a.rank - 1;
            if (q != a.rank && w4 < count4) st_128(a.peer[q] + a.stage_off + static_cast<uint64_t>(slot) * per + v * 4, x[q]);
        }
    }
    // the < 4 trailing words of the whole buffer live in one slice
    for (uint64_t t = count4 * 4 + gtid; t < a.count; t += gstride) {
        const int q = static_cast<int>(t / per);
        const int slot = a.rank < q ? a.rank : a.rank - 1;
        if (q != a.rank) a.peer[q][a.stage_off + static_cast<uint64_t>(slot) * per + (t - static_cast<uint64_t>(q) * per)] = ld_cg32(a.local + t);
    }
}

// Phase B: my slice = my own words + the NR-1 staged copies (all local reads), stored into every rank's buffer.
template <int NR>
__device__ __forceinline__ void push_add_broadcast(const NvlsArgs &a, uint64_t per, uint64_t lo, uint64_t hi, uint64_t gtid,
                                                   uint64_t gstride) {
    const uint64_t lo4 = lo / 4, hi4 = hi / 4;
    const uint32_t *stage = a.local + a.stage_off;
    for (uint64_t v = lo4 + gtid; v < hi4; v += gstride) {
        uint4 x[NR];
        x[0] = ld_cg128(a.local + v * 4);
#pragma unroll
        for (int s = 0; s < NR - 1; ++s) x[s + 1] = ld_cg128(stage + static_cast<uint64_t>(s) * per + (v - lo4) * 4);
        uint4 acc = x[0];
#pragma unroll
        for (int r = 1; r < NR; ++r) {
            acc.x += x[r].x;
            acc.y += x[r].y;
            acc.z += x[r].z;
            acc.w += x[r].w;
        }
#pragma unroll
        for (int r = 0; r < NR; ++r) st_128(a.peer[r] + v * 4, acc);
    }
    for (uint64_t i = max(hi4 * 4, lo) + gtid; i < hi; i += gstride) {
        uint32_t acc = ld_cg32(a.local + i);
        for (int s = 0; s < NR - 1; ++s) acc += ld_cg32(stage + static_cast<uint64_t>(s) * per + (i - lo));
        for (int r = 0; r < NR; ++r) a.peer[r][i] = acc;
    }
}

// All CTAs of all ranks meet here. Local arrival is the cooperative grid barrier; one thread then adds 1 to the
// flag word of EVERY rank through the multicast address (release at system scope, so everything this grid wrote
// before is visible to whoever acquires the flag) and waits until its own copy shows that all ranks did the same.
__device__ void cross_gpu_barrier(cg::grid_group &grid, const NvlsArgs &a, uint32_t index) {
    grid.sync();
    if (blockIdx.x == 0 && threadIdx.x == 0 && !a.barrier_peer) {
        __threadfence_system();
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(a.flag_mc), "r"(1u) : "memory");
        const uint32_t target = (a.barrier_base + index + 1u) * static_cast<uint32_t>(a.nranks);
        const unsigned long long t0 = globaltimer_ns();
        for (;;) {
            uint32_t seen;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.flag_local) : "memory");
            if (static_cast<int32_t>(seen - target) >= 0) break;
            if (globaltimer_ns() - t0 > NVLS_SPIN_TIMEOUT_NS) {
                *a.error = 1;
                break;
            }
            __nanosleep(200);
        }
        __threadfence_system();
    }
    if (blockIdx.x == 0 && threadIdx.x < a.nranks && a.barrier_peer) {
        // No multicast: thread r publishes this rank's epoch into slot [rank] of rank r's flag page with a peer
        // store, then waits until rank r's epoch shows up in slot [r] of the local page.
        const uint64_t flag_off = static_cast<uint64_t>(a.flag_local - a.local);
        const uint32_t epoch = a.barrier_base + index + 1u;
        const int r = threadIdx.x;
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.peer[r] + flag_off + 64 + a.rank), "r"(epoch) : "memory");
        const unsigned long long t0 = globaltimer_ns();
        for (;;) {
            uint32_t seen;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.flag_local + 64 + r) : "memory");
            if (static_cast<int32_t>(seen - epoch) >= 0) break;
            if (globaltimer_ns() - t0 > NVLS_SPIN_TIMEOUT_NS) {
                *a.error = 1;
                break;
            }
            __nanosleep(200);
        }
        __threadfence_system();
    }
    grid.sync();
}

template <int NVLS_UNROLL>
__global__ void __launch_bounds__(NVLS_THREADS) nvls_allreduce_kernel(NvlsArgs a) {
    cg::grid_group grid = cg::this_grid();
    const uint64_t gtid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t gstride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    uint32_t bar = 0;

    if (a.phases & NVLS_FILL) {
        const uint32_t base = hash_base(a.seed, stream_ar(a.rank));
        const uint64_t groups = a.count / 4;
        for (uint64_t g = gtid; g < groups; g += gstride) {
            const uint64_t j = g * 4;
            stg128_stream(a.local + j, make_uint4(hash_at(base, j), hash_at(base, j + 1), hash_at(base, j + 2), hash_at(base, j + 3)));
        }
        for (uint64_t j = groups * 4 + gtid; j < a.count; j += gstride) a.local[j] = hash_at(base, j);
    }

    if ((a.phases & NVLS_REDUCE) && a.algo == 3) {
        // All-push two-shot. No start barrier across GPUs: phase A only reads this rank's own buffer (complete once
        // the local grid has finished filling; peers' broadcast stores of the previous round landed before that round's
        // end barrier) and writes staging slots their owners read only in phase B, after the middle barrier.
        uint64_t per = (a.count + a.nranks - 1) / a.nranks;
        per = (per + 31) & ~31ull;
        const uint64_t lo = min(per * a.rank, a.count), hi = min(lo + per, a.count);
        grid.sync();
        switch (a.nranks) {
        case 2: push_scatter<2>(a, per, gtid, gstride); break;
        case 4: push_scatter<4>(a, per, gtid, gstride); break;
        case 8: push_scatter<8>(a, per, gtid, gstride); break;
        default: break;  // rejected on the host
        }
        cross_gpu_barrier(grid, a, bar++);  // every copy of every slice has reached its owner
        switch (a.nranks) {
        case 2: push_add_broadcast<2>(a, per, lo, hi, gtid, gstride); break;
        case 4: push_add_broadcast<4>(a, per, lo, hi, gtid, gstride); break;
        case 8: push_add_broadcast<8>(a, per, lo, hi, gtid, gstride); break;
        default: break;
        }
        cross_gpu_barrier(grid, a, bar++);  // every slice has been broadcast
    } else if (a.phases & NVLS_REDUCE) {
        cross_gpu_barrier(grid, a, bar++);  // every rank's inputs are in place (and last round's stores have landed)
        // my slice, rounded so that warps stay on 128-byte lines
        uint64_t per = (a.reduce_count + a.nranks - 1) / a.nranks;
        per = (per + 31) & ~31ull;
        const uint64_t lo = min(per * a.rank, a.reduce_count), hi = min(lo + per, a.reduce_count);
        auto p2p_part = [&](uint64_t from, uint64_t to, uint64_t tid0, uint64_t stride0) {
            switch (a.nranks) {
            case 2: p2p_reduce_broadcast<2>(a, from, to, tid0, stride0); break;
            case 4: p2p_reduce_broadcast<4>(a, from, to, tid0, stride0); break;
            case 8: p2p_reduce_broadcast<8>(a, from, to, tid0, stride0); break;
            default: break;  // rejected on the host
            }
        };
        // words [from, to) through the switch: CTA `bid` of `nb` CTAs
        auto nvls_part = [&](uint64_t from, uint64_t to, uint64_t bid, uint64_t nb) {
            constexpr uint64_t TILE = static_cast<uint64_t>(NVLS_THREADS) * NVLS_UNROLL;
            const uint64_t n_tiles = (to - from) / TILE;
            if constexpr (NVLS_UNROLL >= 4) {
                if (a.st_v4) {
                    // Same scalar reduce-loads, but each warp owns NVLS_UNROLL*32 consecutive words of the tile, so four
                    // loads cover 128 consecutive words; a shuffle transpose hands lane l the words 4l..4l+3 of that block
                    // and the broadcast is one 16-byte multimem.st per lane (a quarter of the store instructions).
                    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
                    for (uint64_t t = bid; t < n_tiles; t += nb) {
                        const uint64_t base = from + t * TILE + static_cast<uint64_t>(warp) * (NVLS_UNROLL * 32);
                        uint32_t v[NVLS_UNROLL];
#pragma unroll
                        for (int u = 0; u < NVLS_UNROLL; ++u) v[u] = mm_ld_reduce_add_u32(a.mc + base + u * 32 + lane);
#pragma unroll
                        for (int g = 0; g < NVLS_UNROLL / 4; ++g) {
                            uint32_t o[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int src = (4 * lane + j) & 31;  // word 4*lane+j of the block sits in lane `src` of load lane>>3
                                const uint32_t t0 = __shfl_sync(0xffffffffu, v[4 * g + 0], src), t1 = __shfl_sync(0xffffffffu, v[4 * g + 1], src);
                                const uint32_t t2 = __shfl_sync(0xffffffffu, v[4 * g + 2], src), t3 = __shfl_sync(0xffffffffu, v[4 * g + 3], src);
                                const int sel = lane >> 3;
                                o[j] = sel == 0 ? t0 : sel == 1 ? t1 : sel == 2 ? t2 : t3;
                            }
                            mm_st_v4(a.mc + base + g * 128 + lane * 4, o[0], o[1], o[2], o[3]);
                        }
                    }
                    for (uint64_t i = from + n_tiles * TILE + bid * NVLS_THREADS + threadIdx.x; i < to; i += nb * NVLS_THREADS)
                        mm_st_u32(a.mc + i, mm_ld_reduce_add_u32(a.mc + i));
                    return;
                }
            }
            for (uint64_t t = bid; t < n_tiles; t += nb) {
                const uint64_t base = from + t * TILE + threadIdx.x;
                uint32_t v[NVLS_UNROLL];
#pragma unroll
                for (int u = 0; u < NVLS_UNROLL; ++u) v[u] = mm_ld_reduce_add_u32(a.mc + base + u * NVLS_THREADS);
#pragma unroll
                for (int u = 0; u < NVLS_UNROLL; ++u) mm_st_u32(a.mc + base + u * NVLS_THREADS, v[u]);
            }
            for (uint64_t i = from + n_tiles * TILE + bid * NVLS_THREADS + threadIdx.x; i < to; i += nb * NVLS_THREADS)
                mm_st_u32(a.mc + i, mm_ld_reduce_add_u32(a.mc + i));
        };
        if (a.algo == 1) {
            p2p_part(lo, hi, gtid, gstride);
        } else if (a.algo == 2 && a.hybrid_nvls_blocks > 0 && a.hybrid_nvls_blocks < static_cast<int>(gridDim.x)) {
            // hybrid: the head of the slice goes through the switch, the tail over peer mappings, on disjoint CTAs,
            // so the switch's reduction engine and the spare link bandwidth work at the same time
            const uint64_t nb = static_cast<uint64_t>(a.hybrid_nvls_blocks);
            const uint64_t head = ((hi - lo) / 1000u * static_cast<uint64_t>(a.hybrid_permille)) & ~static_cast<uint64_t>(31);
            const uint64_t mid = (lo + head < hi) ? lo + head : hi;
            if (blockIdx.x < nb)
                nvls_part(lo, mid, blockIdx.x, nb);
            else
                p2p_part(mid, hi, (blockIdx.x - nb) * NVLS_THREADS + threadIdx.x, (gridDim.x - nb) * NVLS_THREADS);
        } else {
            nvls_part(lo, hi, blockIdx.x, gridDim.x);
        }
        cross_gpu_barrier(grid, a, bar++);  // every slice has been broadcast
    }

    if (a.phases & NVLS_VERIFY) {
        unsigned long long bad = 0;
        for (uint64_t i = gtid; i < a.count; i += gstride) {
            uint32_t s = 0;
            for (int r = 0; r < a.nranks; ++r) s += hash(a.seed, stream_ar(r), i);
            uint32_t got;
            asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(got) : "l"(a.local + i) : "memory");  // written by peers: skip L1
            bad += (s * a.mult != got);
        }
        bad = warp_sum_u64(bad);
        if ((threadIdx.x & 31) == 0 && bad) atomicAdd(a.mismatches, bad);
    }
}

// ---- driver entry points, resolved at run time -------------------------------------------------------------
struct DriverApi {
    bool ok = false;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
};
static DriverApi g_drv;
static std::mutex g_drv_mu;

static int load_driver(char *err, size_t errlen) {
    std::lock_guard<std::mutex> lock(g_drv_mu);
    if (g_drv.ok) return GPUPROBE_OK;
    DriverApi d;
#define GP_DRV(field, name)                                                                              \
    {                                                                                                    \
        void *fn = nullptr;                                                                              \
        cudaDriverEntryPointQueryResult q;                                                               \
        cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);                       \
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {                               \
            (void)cudaGetLastError();                                                                    \
            set_err(err, errlen, "driver entry point %s unavailable", name);                             \
            return GPUPROBE_ECUDA;                                                                       \
        }                                                                                                \
        d.field = reinterpret_cast<decltype(d.field)>(fn);                                               \
    }
    GP_DRV(GetErrorString, "cuGetErrorString")
    GP_DRV(DeviceGetAttribute, "cuDeviceGetAttribute")
    GP_DRV(DeviceGet, "cuDeviceGet")
    GP_DRV(MemCreate, "cuMemCreate")
    GP_DRV(MemRelease, "cuMemRelease")
    GP_DRV(MemAddressReserve, "cuMemAddressReserve")
    GP_DRV(MemAddressFree, "cuMemAddressFree")
    GP_DRV(MemMap, "cuMemMap")
    GP_DRV(MemUnmap, "cuMemUnmap")
    GP_DRV(MemSetAccess, "cuMemSetAccess")
    GP_DRV(MemGetAllocationGranularity, "cuMemGetAllocationGranularity")
    GP_DRV(MulticastCreate, "cuMulticastCreate")
    GP_DRV(MulticastAddDevice, "cuMulticastAddDevice")
    GP_DRV(MulticastBindMem, "cuMulticastBindMem")
    GP_DRV(MulticastUnbind, "cuMulticastUnbind")
    GP_DRV(MulticastGetGranularity, "cuMulticastGetGranularity")
#undef GP_DRV
    d.ok = true;
    g_drv = d;
    return GPUPROBE_OK;
}

struct NvlsDev {
    int ord = -1;
    DeviceState *st = nullptr;
    CUmemGenericAllocationHandle mem = 0;
    CUdeviceptr uc = 0;  // unicast mapping
    bool uc_mapped = false, bound = false;
    unsigned long long *bad = nullptr;
    int *error = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int grid = 0;
    // NVLS + copy engines: the stream the engine's round is ordered on, and the two events that tie it to the main stream
    cudaStream_t ce_main = nullptr;
    cudaEvent_t ev_round = nullptr, ev_ce_done = nullptr;
};

}  // namespace gp

using namespace gp;

static int allreduce_symmetric(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, int algo, char *err, size_t errlen) {
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
    if (algo >= 1 && algo <= 3 && nd != 2 && nd != 4 && nd != 8) {
        set_err(err, errlen, "the peer two-shot allreduce is built for 2, 4 or 8 ranks (got %d)", nd);
        return GPUPROBE_EINVAL;
    }
    const char *algo_name = algo == 0 ? "nvls" : algo == 1 ? "two-shot" : algo == 2 ? "hybrid" : algo == 3 ? "push two-shot" : "nvls + copy engines";
    std::vector<NvlsDev> D(nd);
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
    rc = load_driver(err, errlen);
    if (rc) return rc;
    // NVLS needs multicast; the two-shot kernel only needs peer mappings (its barrier then goes through peer stores).
    bool use_mc = !(getenv("GPUPROBE_NO_MULTICAST") && *getenv("GPUPROBE_NO_MULTICAST") == '1');
    for (auto &d : D) {
        int mc = 0;
        if (g_drv.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d.ord) != CUDA_SUCCESS || !mc) use_mc = false;
    }
    if (!use_mc && algo != 1 && algo != 3) {
        set_err(err, errlen, "NVLink multicast (NVLS) is not available on every device; use gpuprobe_allreduce_i32_p2p");
        return GPUPROBE_ENOPEER;
    }
    if (!use_mc) {
        for (auto &x : D)
            for (auto &y : D) {
                int can = 1;
                if (x.ord != y.ord && (cudaDeviceCanAccessPeer(&can, x.ord, y.ord) != cudaSuccess || !can)) {
                    (void)cudaGetLastError();
                    set_err(err, errlen, "device %d cannot access device %d", x.ord, y.ord);
                    return GPUPROBE_ENOPEER;
                }
            }
    }
    std::vector<int> order(nd);
    for (int i = 0; i < nd; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return D[x].ord < D[y].ord; });
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int i : order) locks.emplace_back(D[i].st->mu);

    // tuning knobs (defaults are the measured best; see profiles/)
    // algo 4 keeps half of every SM free: the copy-engine side's add kernel has to run beside the resident multimem kernel
    int unroll = 8, ctas_per_sm = algo == 4 ? 1 : 2;
    if (const char *u = getenv("GPUPROBE_NVLS_UNROLL")) unroll = atoi(u);
    if (const char *c = getenv("GPUPROBE_NVLS_CTAS")) ctas_per_sm = std::max(1, atoi(c));
    int st_v4 = 0;  // 16-byte multimem.st: measured in profiles/r02_fabric.md
    if (const char *v = getenv("GPUPROBE_NVLS_ST_V4")) st_v4 = atoi(v) ? 1 : 0;
    int hybrid_permille = 950, hybrid_block_permille = 350;  // centre of the measured plateau at 8 ranks (profiles/r01_fabric.md)
    if (const char *h = getenv("GPUPROBE_HYBRID_PERMILLE")) hybrid_permille = std::max(0, std::min(1000, atoi(h)));
    if (const char *h = getenv("GPUPROBE_HYBRID_BLOCK_PERMILLE")) hybrid_block_permille = std::max(1, std::min(999, atoi(h)));
    const void *kernel_fn = unroll >= 16  ? reinterpret_cast<const void *>(nvls_allreduce_kernel<16>)
                            : unroll >= 8 ? reinterpret_cast<const void *>(nvls_allreduce_kernel<8>)
                            : unroll >= 4 ? reinterpret_cast<const void *>(nvls_allreduce_kernel<4>)
                                          : reinterpret_cast<const void *>(nvls_allreduce_kernel<2>);

    const uint64_t payload = cfg->count * sizeof(uint32_t);
    // all-push two-shot: behind the payload every rank keeps nd-1 staging slices for its peers' copies of its slice
    uint64_t per_words = (cfg->count + nd - 1) / nd;
    per_words = (per_words + 31) & ~31ull;  // as in the kernel
    const uint64_t stage_off_words = (cfg->count + 31) & ~31ull;
    // NVLS + copy engines (algo 4): the multimem kernel reduces words [0, head), the copy-engine two-shot words [head, count)
    // at the same time; the engine's staging slices sit behind the payload like the push kernel's
    int nvlsce_permille = 800;
    if (const char *h = getenv("GPUPROBE_NVLSCE_PERMILLE")) nvlsce_permille = std::max(1, std::min(999, atoi(h)));
    uint64_t head_words = cfg->count;
    if (algo == 4) head_words = std::min<uint64_t>(cfg->count, (cfg->count / 1000u * static_cast<uint64_t>(nvlsce_permille) + 31) & ~31ull);
    const uint64_t tail_words = cfg->count - head_words;
    const uint64_t data_bytes = algo == 3   ? (stage_off_words + static_cast<uint64_t>(nd - 1) * per_words) * sizeof(uint32_t)
                                : algo == 4 ? (stage_off_words + CeTwoShot::stage_words(tail_words, nd)) * sizeof(uint32_t)
                                            : payload;
    CeTwoShot engine;
    CUmemGenericAllocationHandle mc_handle = 0;
    CUdeviceptr mc_va = 0;
    bool mc_created = false, mc_mapped = false;
    size_t size = 0;

    auto cleanup = [&]() {
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaDeviceSynchronize();
        }
        engine.destroy();
        for (auto &d : D) {
            if (cudaSetDevice(d.ord) != cudaSuccess) continue;
            cudaFree(d.bad);
            cudaFree(d.error);
            if (d.ev0) cudaEventDestroy(d.ev0);
            if (d.ev1) cudaEventDestroy(d.ev1);
            if (d.ev_round) cudaEventDestroy(d.ev_round);
            if (d.ev_ce_done) cudaEventDestroy(d.ev_ce_done);
            if (d.ce_main) cudaStreamDestroy(d.ce_main);
        }
        if (mc_mapped) g_drv.MemUnmap(mc_va, size);
        if (mc_va) g_drv.MemAddressFree(mc_va, size);
        for (auto &d : D) {
            if (d.bound) g_drv.MulticastUnbind(mc_handle, d.ord, 0, size);
            if (d.uc_mapped) g_drv.MemUnmap(d.uc, size);
            if (d.uc) g_drv.MemAddressFree(d.uc, size);
            if (d.mem) g_drv.MemRelease(d.mem);
        }
        if (mc_created) g_drv.MemRelease(mc_handle);
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
#define GP_DRVTRY(call)                                                                                \
    do {                                                                                               \
        CUresult r_ = (call);                                                                          \
        if (r_ != CUDA_SUCCESS) {                                                                      \
            const char *s_ = "?";                                                                      \
            g_drv.GetErrorString(r_, &s_);                                                             \
            set_err(err, errlen, "%s failed: %s (%s:%d)", #call, s_, __FILE__, __LINE__);              \
            cleanup();                                                                                 \
            return r_ == CUDA_ERROR_OUT_OF_MEMORY ? GPUPROBE_ENOMEM : GPUPROBE_ECUDA;                  \
        }                                                                                              \
    } while (0)

    // one multicast object over all ranks; every rank binds a physical allocation of the same size
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(nd);
    mp.handleTypes = CU_MEM_HANDLE_TYPE_NONE;
    mp.size = data_bytes + NVLS_FLAG_BYTES;
    size_t gran = 0;
    if (use_mc) GP_DRVTRY(g_drv.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    for (auto &d : D) {
        CUmemAllocationProp ap;
        std::memset(&ap, 0, sizeof(ap));
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = d.ord;
        size_t g2 = 0;
        GP_DRVTRY(g_drv.MemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
        gran = std::max(gran, g2);
    }
    size = ((data_bytes + NVLS_FLAG_BYTES + gran - 1) / gran) * gran;
    mp.size = size;
    const uint64_t flag_off = size - NVLS_FLAG_BYTES;  // last 4 KiB of the allocation, past the payload
    if (use_mc) {
        GP_DRVTRY(g_drv.MulticastCreate(&mc_handle, &mp));
        mc_created = true;
        for (auto &d : D) GP_DRVTRY(g_drv.MulticastAddDevice(mc_handle, d.ord));
    }
    std::vector<CUmemAccessDesc> access(nd);
    for (int i = 0; i < nd; ++i) {
        access[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        access[i].location.id = D[i].ord;
        access[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    for (int i = 0; i < nd; ++i) {
        NvlsDev &d = D[i];
        GP_TRY(cudaSetDevice(d.ord));
        CUmemAllocationProp ap;
        std::memset(&ap, 0, sizeof(ap));
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id = d.ord;
        GP_DRVTRY(g_drv.MemCreate(&d.mem, size, &ap, 0));
        GP_DRVTRY(g_drv.MemAddressReserve(&d.uc, size, gran, 0, 0));
        GP_DRVTRY(g_drv.MemMap(d.uc, size, 0, d.mem, 0));
        d.uc_mapped = true;
        // every rank may read and write every rank's buffer (the two-shot algorithm; harmless for NVLS)
        GP_DRVTRY(g_drv.MemSetAccess(d.uc, size, access.data(), static_cast<size_t>(nd)));
        if (use_mc) {
            GP_DRVTRY(g_drv.MulticastBindMem(mc_handle, 0, d.mem, 0, size, 0));
            d.bound = true;
        }
        GP_TRY(cudaMalloc(&d.bad, sizeof(unsigned long long)));
        GP_TRY(cudaMalloc(&d.error, sizeof(int)));
        GP_TRY(cudaMemsetAsync(d.bad, 0, sizeof(unsigned long long), d.st->stream));
        GP_TRY(cudaMemsetAsync(d.error, 0, sizeof(int), d.st->stream));
        GP_TRY(cudaMemsetAsync(reinterpret_cast<void *>(d.uc + flag_off), 0, NVLS_FLAG_BYTES, d.st->stream));
        GP_TRY(cudaEventCreate(&d.ev0));
        GP_TRY(cudaEventCreate(&d.ev1));
        int per_sm = 0;
        GP_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel_fn, NVLS_THREADS, 0));
        d.grid = d.st->sm_count * std::max(1, std::min(per_sm, ctas_per_sm));
    }
    if (use_mc) {
        GP_DRVTRY(g_drv.MemAddressReserve(&mc_va, size, gran, 0, 0));
        GP_DRVTRY(g_drv.MemMap(mc_va, size, 0, mc_handle, 0));
        mc_mapped = true;
        GP_DRVTRY(g_drv.MemSetAccess(mc_va, size, access.data(), static_cast<size_t>(nd)));
    }
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        GP_TRY(cudaStreamSynchronize(d.st->stream));
    }

    const bool ce_tail = algo == 4 && tail_words > 0;
    if (ce_tail) {
        std::vector<CeRank> ranks(nd);
        for (int r = 0; r < nd; ++r) {
            NvlsDev &d = D[r];
            GP_TRY(cudaSetDevice(d.ord));
            GP_TRY(cudaStreamCreateWithFlags(&d.ce_main, cudaStreamNonBlocking));
            GP_TRY(cudaEventCreateWithFlags(&d.ev_round, cudaEventDisableTiming));
            GP_TRY(cudaEventCreateWithFlags(&d.ev_ce_done, cudaEventDisableTiming));
            ranks[r].ord = d.ord;
            ranks[r].sm_count = d.st->sm_count;
            ranks[r].buf = reinterpret_cast<uint32_t *>(d.uc) + head_words;
            ranks[r].stage = reinterpret_cast<uint32_t *>(d.uc) + stage_off_words;
            ranks[r].main = d.ce_main;
        }
        GP_TRY(engine.init(ranks, tail_words, 0));
    }

    uint32_t barriers_done = 0;
    auto launch_all = [&](int phases, uint32_t mult, bool timed) -> cudaError_t {
        for (int r = 0; r < nd; ++r) {
            NvlsDev &d = D[r];
            cudaError_t e = cudaSetDevice(d.ord);
            if (e != cudaSuccess) return e;
            NvlsArgs a;
            a.local = reinterpret_cast<uint32_t *>(d.uc);
            a.mc = reinterpret_cast<uint32_t *>(mc_va);
            a.flag_local = reinterpret_cast<uint32_t *>(d.uc + flag_off);
            a.flag_mc = reinterpret_cast<uint32_t *>(mc_va + flag_off);
            a.mismatches = d.bad;
            a.error = d.error;
            a.count = cfg->count;
            a.reduce_count = head_words;
            a.seed = cfg->seed;
            a.mult = mult;
            a.barrier_base = barriers_done;
            a.rank = r;
            a.nranks = nd;
            a.phases = phases;
            a.algo = algo == 4 ? 0 : algo;
            a.barrier_peer = use_mc ? 0 : 1;
            a.st_v4 = st_v4;
            a.stage_off = stage_off_words;
            a.hybrid_permille = hybrid_permille;
            a.hybrid_nvls_blocks = std::max(1, std::min(d.grid - 1, d.grid * hybrid_block_permille / 1000));
            for (int q = 0; q < NVLS_MAX_PEERS; ++q) a.peer[q] = q < nd ? reinterpret_cast<uint32_t *>(D[q].uc) : nullptr;
            void *args[] = {&a};
            if (timed && (e = cudaEventRecord(d.ev0, d.st->stream)) != cudaSuccess) return e;
            if (ce_tail && (phases & NVLS_REDUCE)) {  // the engine's round starts with the kernel, not after it
                if ((e = cudaEventRecord(d.ev_round, d.st->stream)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(d.ce_main, d.ev_round, 0)) != cudaSuccess) return e;
            }
            e = cudaLaunchCooperativeKernel(kernel_fn, dim3(d.grid), dim3(NVLS_THREADS),
                                            args, 0, d.st->stream);
            if (e != cudaSuccess) return e;
        }
        if (ce_tail && (phases & NVLS_REDUCE)) {
            cudaError_t e = engine.issue_round();
            if (e != cudaSuccess) return e;
            for (auto &d : D) {  // the round is over on a rank when its kernel and the engine's round both are
                if ((e = cudaSetDevice(d.ord)) != cudaSuccess) return e;
                if ((e = cudaEventRecord(d.ev_ce_done, d.ce_main)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(d.st->stream, d.ev_ce_done, 0)) != cudaSuccess) return e;
            }
        }
        if (timed)
            for (auto &d : D) {
                cudaError_t e = cudaSetDevice(d.ord);
                if (e == cudaSuccess) e = cudaEventRecord(d.ev1, d.st->stream);
                if (e != cudaSuccess) return e;
            }
        if (phases & NVLS_REDUCE) barriers_done += 2;
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

    // round 1 is the fused launch: fill -> barrier -> reduce+broadcast -> barrier (no verify yet). With the copy engines
    // in the round the fill is a launch of its own: the engines read the tail as soon as the round starts.
    if (ce_tail) {
        GP_TRY(launch_all(NVLS_FILL, 1, false));
        GP_TRY(sync_all());
        GP_TRY(launch_all(NVLS_REDUCE, 1, false));
    } else {
        GP_TRY(launch_all(NVLS_FILL | NVLS_REDUCE, 1, false));
    }
    GP_TRY(sync_all());
    int rounds = 1;
    for (int i = 1; i < cfg->warmup; ++i, ++rounds) {
        GP_TRY(launch_all(NVLS_REDUCE, 1, false));
        GP_TRY(sync_all());
    }
    std::vector<float> ms;
    for (int i = 0; i < cfg->iters; ++i, ++rounds) {
        GP_TRY(launch_all(NVLS_REDUCE, 1, true));
        GP_TRY(sync_all());
        float worst = 0;
        for (auto &d : D) {
            float m = 0;
            GP_TRY(cudaEventElapsedTime(&m, d.ev0, d.ev1));
            worst = std::max(worst, m);
        }
        ms.push_back(worst);
    }
    // after `rounds` in-place rounds every word is n^(rounds-1) * sum_r x_r[i]  (mod 2^32)
    uint32_t mult = 1;
    for (int k = 1; k < rounds; ++k) mult *= static_cast<uint32_t>(nd);
    GP_TRY(launch_all(NVLS_VERIFY, mult, false));
    GP_TRY(sync_all());
    uint64_t bad_total = 0;
    int timed_out = 0;
    for (auto &d : D) {
        GP_TRY(cudaSetDevice(d.ord));
        unsigned long long h = 0;
        int e = 0;
        GP_TRY(cudaMemcpy(&h, d.bad, sizeof(h), cudaMemcpyDeviceToHost));
        GP_TRY(cudaMemcpy(&e, d.error, sizeof(e), cudaMemcpyDeviceToHost));
        bad_total += h;
        timed_out |= e;
    }
    // What the ranks hold now, digested by the same kernel that digests vectorAdd's C: the caller compares it with a
    // value computed elsewhere (tests: the CPU oracle) -- `mismatches` above comes from this library's own generator.
    out->ranks_equal = 1;
    for (int r = 0; r < nd; ++r) {
        NvlsDev &d = D[r];
        GP_TRY(cudaSetDevice(d.ord));
        uint64_t s64 = 0;
        uint32_t x32 = 0;
        rc = run_digest(*d.st, reinterpret_cast<const void *>(d.uc), cfg->count, &s64, &x32, d.st->stream, err, errlen);
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
    out->rounds = static_cast<uint32_t>(rounds);
    out->impl = algo == 0 ? GPUPROBE_AR_IMPL_NVLS : algo == 1 ? GPUPROBE_AR_IMPL_P2P : algo == 2 ? GPUPROBE_AR_IMPL_HYBRID
                : algo == 3 ? GPUPROBE_AR_IMPL_PUSH : GPUPROBE_AR_IMPL_NVLS_CE;
    out->ndev = nd;
    out->nccl_version = 0;  // not NCCL
    out->bytes = payload;
    out->mismatches = bad_total;
    out->ms_best = *std::min_element(ms.begin(), ms.end());
    out->ms_median = median_of(ms);
    out->algbw_gbps = out->ms_median > 0 ? payload / (out->ms_median * 1e6) : 0;
    out->busbw_gbps = out->algbw_gbps * 2.0 * (nd - 1) / nd;
    out->frac_of_peak = out->busbw_gbps / (cfg->peak_gbps > 0 ? cfg->peak_gbps : 900.0);
    cleanup();
    if (timed_out) {
        set_err(err, errlen, "allreduce kernel (%s): a cross-GPU barrier timed out", algo_name);
        return GPUPROBE_ECUDA;
    }
    if (bad_total) {
        set_err(err, errlen, "allreduce kernel (%s): %llu elements differ from the expected sum", algo_name,
                (unsigned long long)bad_total);
        return GPUPROBE_FAIL_VERIFY;
    }
    if (!out->ranks_equal) {
        set_err(err, errlen, "allreduce kernel (%s): the ranks' buffers differ from one another", algo_name);
        return GPUPROBE_FAIL_VERIFY;
    }
    return GPUPROBE_OK;
#undef GP_TRY
#undef GP_DRVTRY
}

extern "C" int gpuprobe_allreduce_i32_nvls(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    return allreduce_symmetric(cfg, out, 0, err, errlen);
}

extern "C" int gpuprobe_allreduce_i32_hybrid(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    return allreduce_symmetric(cfg, out, 2, err, errlen);
}

extern "C" int gpuprobe_allreduce_i32_p2p(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    return allreduce_symmetric(cfg, out, 1, err, errlen);
}

extern "C" int gpuprobe_allreduce_i32_nvls_ce(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    return allreduce_symmetric(cfg, out, 4, err, errlen);
}

extern "C" int gpuprobe_allreduce_i32_push(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    return allreduce_symmetric(cfg, out, 3, err, errlen);
}

// The kernel the validator runs when it is not told which: the measured best per rank count
// (profiles/r02_fabric.md). Falls back to the peer-mapping kernels where NVLS is unavailable.
extern "C" int gpuprobe_allreduce_i32_auto(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen) {
    if (!cfg || !out) {
        set_err(err, errlen, "null cfg/out");
        return GPUPROBE_EINVAL;
    }
    const int nd = cfg->ndev;
    const uint64_t bytes = cfg->count * sizeof(uint32_t);
    // Measured (profiles/r02_fabric.md): large buffers at 2 and 4 ranks are fastest with the copy engines moving the data
    // (748 / 717 GB/s busbw at 1 GiB against 646 / 684 for the best kernel); below the crossover the engines' start-up
    // latency loses to one cooperative kernel; from 5 ranks up the engines' many concurrent copies lose to the switch
    // (8 ranks: 426 against 816). Other rank counts than 2, 4, 8: NVLS, or the engines at 3 ranks.
    int algo = nd == 2 ? 3 : 2;  // 2 ranks: all-push two-shot; 4 and 8: hybrid (815 GB/s at 8)
    if (nd != 2 && nd != 4 && nd != 8) algo = 0;
    const bool engines = (nd == 2 && bytes >= (32ull << 20)) || ((nd == 3 || nd == 4) && bytes >= (512ull << 20));
    if (const char *e = getenv("GPUPROBE_AR_AUTO_ALGO")) {
        algo = std::max(0, std::min(5, atoi(e)));
        if (algo == 5) return gpuprobe_allreduce_i32_ce(cfg, out, err, errlen);
    } else if (engines) {
        return gpuprobe_allreduce_i32_ce(cfg, out, err, errlen);
    }
    int rc = allreduce_symmetric(cfg, out, algo, err, errlen);
    if (rc == GPUPROBE_ENOPEER && algo != 1)  // no multicast on this box: peer mappings (2, 4, 8 ranks) or the copy engines
        rc = (nd == 2 || nd == 4 || nd == 8) ? allreduce_symmetric(cfg, out, 1, err, errlen) : gpuprobe_allreduce_i32_ce(cfg, out, err, errlen);
    return rc;
}

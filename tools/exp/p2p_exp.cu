// p2p_exp.cu -- scratch experiments behind the P2P kernel's shape (not part of the product).
// GPU 1 pulls 1 GiB from GPU 0 (peer loads), stores locally and checksums; variants: vector width, loads in flight, CTAs per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/exp/bin/p2p_exp tools/exp/p2p_exp.cu
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(2); } } while (0)

template <int VB> struct V;
template <> struct V<16> { uint32_t w[4]; };
template <> struct V<32> { uint32_t w[8]; };

template <int VB, int HINT> __device__ __forceinline__ V<VB> ldv(const void *p) {
    V<VB> r;
    if constexpr (VB == 16) {
        if constexpr (HINT == 0) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
        else asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]) : "l"(p));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]) : "l"(p));
    }
    return r;
}
template <int VB> __device__ __forceinline__ void stv(void *p, const V<VB> &r) {
    if constexpr (VB == 16) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]) : "memory");
    else asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]), "r"(r.w[4]), "r"(r.w[5]), "r"(r.w[6]), "r"(r.w[7]) : "memory");
}

template <int VB, int U, int T, int HINT>
__global__ void __launch_bounds__(T) k_pull(const char *__restrict__ src, char *__restrict__ dst, size_t nvec, unsigned long long *sum) {
    unsigned long long s = 0;
    const size_t tile = (size_t)T * U, n_tiles = nvec / tile;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        V<VB> v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldv<VB, HINT>(src + (base + (size_t)u * T) * VB);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            stv<VB>(dst + (base + (size_t)u * T) * VB, v[u]);
#pragma unroll
            for (int i = 0; i < VB / 4; ++i) s += v[u].w[i];
        }
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(sum, s);
}

int main(int argc, char **argv) {
    const size_t bytes = (size_t)1 << 30;
    int iters = argc > 1 ? atoi(argv[1]) : 10;
    int n; CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    char *src, *dst; unsigned long long *sum;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&src, bytes)); CK(cudaMemset(src, 0x5a, bytes)); CK(cudaDeviceSynchronize());
    CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&dst, bytes)); CK(cudaMalloc(&sum, 8));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 1));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto run = [&](const char *name, auto kern, int vb, int T, int ctas) {
        CK(cudaMemset(sum, 0, 8));
        for (int i = 0; i < 3; ++i) kern<<<sms * ctas, T>>>(src, dst, bytes / vb, sum);
        CK(cudaGetLastError());
        std::vector<float> ms(iters);
        for (int i = 0; i < iters; ++i) {
            CK(cudaEventRecord(e0)); kern<<<sms * ctas, T>>>(src, dst, bytes / vb, sum); CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms[i], e0, e1));
        }
        std::sort(ms.begin(), ms.end());
        printf("{\"exp\": \"%s_x%d\", \"gbps_median\": %.1f, \"gbps_best\": %.1f}\n", name, ctas, bytes / (ms[iters / 2] * 1e6), bytes / (ms[0] * 1e6));
        fflush(stdout);
    };
    for (int ctas : {1, 2, 4}) {
        run("v16_u4_t512", k_pull<16, 4, 512, 0>, 16, 512, ctas);
        run("v16_u8_t512", k_pull<16, 8, 512, 0>, 16, 512, ctas);
        run("v16_u16_t512", k_pull<16, 16, 512, 0>, 16, 512, ctas);
        run("v32_u2_t512", k_pull<32, 2, 512, 0>, 32, 512, ctas);
        run("v32_u4_t512", k_pull<32, 4, 512, 0>, 32, 512, ctas);
        run("v32_u8_t512", k_pull<32, 8, 512, 0>, 32, 512, ctas);
        run("v16_u8_t1024", k_pull<16, 8, 1024, 0>, 16, 1024, ctas);
        run("v16_u8_t256", k_pull<16, 8, 256, 0>, 16, 256, ctas);
        run("v16_u8_t512_relaxedsys", k_pull<16, 8, 512, 1>, 16, 512, ctas);
    }
    return 0;
}

// gen.cuh -- the probe's deterministic input generator (host + device).
//
// Counter-based, so any slice [first, first+n) of any stream is reproducible on the host,
// on the device and in the oracle without materialising the rest (DESIGN.md "Generator").
// It stands in for the sample's rand()/(float)RAND_MAX fill of h_A and h_B (the external
// vectorAdd binary run at reference validator/cuda-workload-validation.yaml:19-20).
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define GP_HD __host__ __device__ __forceinline__
#else
#define GP_HD inline
#endif

namespace gp {

constexpr uint32_t STREAM_A = 0u;
constexpr uint32_t STREAM_B = 1u;
GP_HD uint32_t stream_p2p(int src) { return 0x100u + static_cast<uint32_t>(src); }
GP_HD uint32_t stream_ar(int rank) { return 0x200u + static_cast<uint32_t>(rank); }

GP_HD uint32_t mix32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

// Per-launch constant part of the hash: depends on (seed, stream) only.
GP_HD uint32_t hash_base(uint32_t seed, uint32_t stream) { return mix32(seed ^ ((stream + 1u) * 0x9E3779B9U)); }

GP_HD uint32_t hash_at(uint32_t base, uint64_t i) {
    uint32_t h = mix32(base ^ static_cast<uint32_t>(i >> 32));
    return mix32(h ^ static_cast<uint32_t>(i));
}

GP_HD uint32_t hash(uint32_t seed, uint32_t stream, uint64_t i) { return hash_at(hash_base(seed, stream), i); }

// value bits for element i of stream A (is_b = false) or B (is_b = true); kind = GPUPROBE_GEN_*
GP_HD uint32_t gen_bits(uint32_t base, uint64_t i, int kind, bool is_b) {
    const uint32_t h = hash_at(base, i);
    if (kind == 1) {
        // 16 x 16 IEEE-754 edge values. B holds finite values only, so A + B is never
        // (+inf) + (-inf): NaN payloads legitimately differ between x86 and the GPU.
        const uint32_t k = h & 15u;
        const uint32_t sign = (k & 1u) << 31;
        uint32_t mag;
        if (k < 8u) {  // +-0, +-min denormal, +-max denormal, +-min normal (same in A and B)
            const uint32_t m = k >> 1;
            mag = m == 0u ? 0u : (m == 1u ? 0x00000001u : (m == 2u ? 0x007fffffu : 0x00800000u));
            return sign | mag;
        }
        if (!is_b) {
            switch (k) {
            case 8: return 0x3f7fffffu;   // 1 - 2^-24
            case 9: return 0x3f800001u;   // 1 + 2^-23
            case 10: return 0x3f800000u;  // 1
            case 11: return 0xbf800000u;  // -1
            case 12: return 0x7f7fffffu;  // max finite
            case 13: return 0xff7fffffu;
            case 14: return 0x7f800000u;  // +inf
            default: return 0xff800000u;  // -inf
            }
        }
        switch (k) {
        case 8: return 0x33800000u;   // 2^-24  (half ulp of 1)
        case 9: return 0x33000000u;   // 2^-25
        case 10: return 0xb3800000u;  // -2^-24
        case 11: return 0xbf7fffffu;  // -(1 - 2^-24)
        case 12: return 0xbf800001u;  // -(1 + 2^-23)
        case 13: return 0x7f7fffffu;
        case 14: return 0xff7fffffu;
        default: return 0x73000000u;  // 2^103 (half ulp of max finite)
        }
    }
    if (kind == 2) {
        // arbitrary finite pattern: an all-ones exponent becomes 0xfe
        return (((h >> 23) & 0xffu) == 0xffu) ? (h ^ 0x00800000u) : h;
    }
    // 24-bit integer * 2^-24: exact in binary32, in [0,1)
    const float f = static_cast<float>(h >> 8) * (1.0f / 16777216.0f);
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
#endif
}

}  // namespace gp

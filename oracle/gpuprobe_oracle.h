/*
 * gpuprobe_oracle.h -- CPU oracle for the validator's CUDA-workload path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under gpu-operator_b200/ may include, link
 * or call this. The only users are tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py.
 *
 * PARITY UNPINNED. The arithmetic of this path is not in the reference tree:
 * it is the prebuilt third-party `vectorAdd` sample binary that the validator
 * image copies out of nvcr.io/nvidia/k8s/cuda-sample:vectoradd-cuda11.7.1-<dist>
 * (reference validator/Dockerfile:33-35,49; version pin validator/versions.mk:23
 * CUDA_SAMPLES_VERSION ?= 11.7.1; call sites validator/cuda-workload-validation.yaml:19-20
 * and validator/plugin-workload-validation.yaml:19-20). Its source is absent
 * from /root/reference and the reference holds no golden vector, known-answer
 * test or fixture for it; the only pinned contract is "exit 0 iff C == A + B"
 * (pod phase Succeeded, validator/main.go:1049). This file therefore restates
 * the sample's published algorithm -- C[i] = A[i] + B[i] in IEEE-754 binary32,
 * round-to-nearest-even, followed by a host-side element check -- and is
 * anchored by hand-derived IEEE-754 known answers, by a binary32 adder written
 * from the standard in pure integer arithmetic (oracle/ieee754_add.py, checked
 * against this loop on millions of operand pairs in tests/test_oracle_ieee754.py),
 * plus the fixtures in tests/golden/ that tests/golden/make_golden.py generates
 * from an independent numpy restatement (oracle/oracle_np.py).
 *
 * The input generator, the digests and the fabric-probe expected values are
 * NEW (they have no reference counterpart; see DESIGN.md "Generator"): the
 * oracle restates their specification independently of the CUDA code.
 */
#ifndef GPUPROBE_ORACLE_H
#define GPUPROBE_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* generator kinds (DESIGN.md "Generator") */
#define ORACLE_GEN_UNIT 0 /* u24(hash) * 2^-24, in [0,1): the sample's rand()/RAND_MAX stand-in */
#define ORACLE_GEN_EDGE 1 /* 16x16 table of IEEE-754 edge values (no NaN can result)            */
#define ORACLE_GEN_BITS 2 /* arbitrary finite bit patterns incl. denormals, both signs          */

/* stream ids fed to the hash */
#define ORACLE_STREAM_A 0u
#define ORACLE_STREAM_B 1u
#define ORACLE_STREAM_P2P(src) (0x100u + (uint32_t)(src))
#define ORACLE_STREAM_AR(rank) (0x200u + (uint32_t)(rank))

uint32_t oracle_mix32(uint32_t x);
uint32_t oracle_hash(uint32_t seed, uint32_t stream, uint64_t i);

/* bit pattern of generated element i of stream A or B */
uint32_t oracle_gen_bits(uint32_t seed, uint32_t stream, uint64_t i, int kind);

/* a[j], b[j] for j in [0,n) hold elements first+j */
void oracle_fill_inputs(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int kind);

void oracle_fill_inputs_mt(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int kind, int threads);

/* The host loop: c[i] = a[i] + b[i], fp32, one rounding (the "Go host loop" of north_star,
 * the verification loop of the vectorAdd sample). */
void oracle_vectoradd(const float *a, const float *b, float *c, uint64_t n);

/* same loop split over `threads` pthreads (CPU baseline with all host cores) */
void oracle_vectoradd_mt(const float *a, const float *b, float *c, uint64_t n, int threads);

/* order-independent digests of the bit patterns of c */
void oracle_digest(const float *c, uint64_t n, uint64_t *sum64, uint32_t *xor32);

/* digest of C over elements [first, first+n) without materialising arrays
 * (chunk-composable: sums add mod 2^64, xors xor) */
void oracle_vectoradd_digest(uint64_t first, uint64_t n, uint32_t seed, int kind, int threads,
                             uint64_t *sum64, uint32_t *xor32);

/* sample's tolerance check fabs(a+b-c) > 1e-5, returns number of failing elements;
 * subsumed by bit equality but restated for completeness */
uint64_t oracle_sample_check(const float *a, const float *b, const float *c, uint64_t n);

/* number of elements whose bit patterns differ */
uint64_t oracle_count_mismatch(const float *x, const float *y, uint64_t n);

/* P2P probe: source buffer words and their checksum */
void oracle_p2p_fill(uint32_t *w, uint64_t first, uint64_t n, uint32_t seed, int src);
void oracle_p2p_checksum(uint64_t nwords, uint32_t seed, int src, int threads,
                         uint64_t *sum64, uint32_t *xor32);

/* allreduce probe: rank r contributes (int32)hash(seed, AR(r), i); expected = wraparound sum */
void oracle_allreduce_fill(int32_t *x, uint64_t first, uint64_t n, uint32_t seed, int rank);
void oracle_allreduce_expected(int32_t *out, uint64_t first, uint64_t n, uint32_t seed, int nranks);
/* sum64 / xor32 of the buffer every rank holds after `rounds` in-place allreduces (threaded) */
void oracle_allreduce_digest(uint64_t first, uint64_t n, uint32_t seed, int nranks, int rounds, int threads,
                             uint64_t *sum64, uint32_t *xor32);

/* 1 if the calling thread's FP environment is IEEE (no FTZ/DAZ, round-to-nearest) */
int oracle_fp_env_ok(void);

#ifdef __cplusplus
}
#endif
#endif

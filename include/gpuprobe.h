/*
 * gpuprobe.h -- C ABI of libgpuprobe.so: B200 (sm_100a) node-validation probes for the
 * GPU Operator validator.
 *
 * The reference has no FFI on this path: its boundary is a process boundary -- the
 * validator creates a Pod whose init container runs `sh -c vectorAdd`
 * (reference validator/cuda-workload-validation.yaml:19-20,
 * validator/plugin-workload-validation.yaml:19-20) and only the exit code is read
 * (pod phase Succeeded, validator/main.go:1042-1059).  This header is the library a
 * maintainer would bind from Go with cgo inside that workload container (INTEGRATION.md
 * shows the stub); each entry point cites the reference behaviour it stands in for.
 *
 * Conventions
 *   return 0   probe ran and passed                        (process exit 0)
 *   return >0  probe ran and FAILED verification           (GPUPROBE_FAIL_*; exit 1)
 *   return <0  probe could not run                         (GPUPROBE_E*;     exit 1)
 *   `err` (may be NULL) receives a NUL-terminated message of at most errlen bytes.
 *   Nothing is thrown across the ABI and there are no callbacks.  The caller owns every
 *   struct and buffer it passes.  Unless an entry point says otherwise the library
 *   allocates the device memory it needs: the probes' arrays are freed before the call
 *   returns; two small caches live per device until gpuprobe_shutdown() -- the digest /
 *   verify scratch (about 14 KiB) and the staging slots of gpuprobe_vectoradd_buffers
 *   (as large as the largest chunk staged so far; env GPUPROBE_BUF_CACHE=0 frees them after
 *   every call).  Every entry point selects its device itself (safe from any OS thread,
 *   e.g. a Go goroutine that migrated).  The synchronous probes (gpuprobe_vectoradd,
 *   _vectoradd_buffers, _p2p_matrix, _allreduce_*) are serialised per device inside the
 *   library; the *_device entry points only enqueue work on the caller's stream and
 *   take no probe lock (digest and verify serialise on their shared scratch).
 *   There is no CPU fallback: without a usable sm_100 device the compute entry points
 *   return GPUPROBE_ENODEV / GPUPROBE_EARCH.
 */
#ifndef GPUPROBE_H
#define GPUPROBE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GPUPROBE_API __attribute__((visibility("default")))
#else
#define GPUPROBE_API
#endif

#define GPUPROBE_ABI_VERSION 2 /* 2: gpuprobe_ar_result carries the digest of the reduced buffer */

/* ---- status codes ------------------------------------------------------------- */
#define GPUPROBE_OK 0
#define GPUPROBE_FAIL_VERIFY 1   /* output differs from the expected values             */
#define GPUPROBE_FAIL_BANDWIDTH 2 /* verified, but below the caller's min_frac threshold */
#define GPUPROBE_EINVAL (-1)     /* bad argument                                        */
#define GPUPROBE_ENODEV (-2)     /* no such CUDA device / driver not loadable           */
#define GPUPROBE_EARCH (-3)      /* device is not sm_100 (kernels are sm_100a only)     */
#define GPUPROBE_ECUDA (-4)      /* a CUDA runtime call failed (message in err)         */
#define GPUPROBE_ENOMEM (-5)     /* device or pinned-host allocation failed             */
#define GPUPROBE_ENOPEER (-6)    /* peer access between two devices is unavailable      */
#define GPUPROBE_ENCCL (-7)      /* NCCL could not be loaded or returned an error       */

/* ---- generator kinds (DESIGN.md "Generator") ------------------------------------ */
#define GPUPROBE_GEN_UNIT 0 /* a,b uniform in [0,1): what the sample's rand()/RAND_MAX gives */
#define GPUPROBE_GEN_EDGE 1 /* IEEE-754 edge values: +-0, denormals, +-inf, max finite, ties */
#define GPUPROBE_GEN_BITS 2 /* arbitrary finite bit patterns                                  */

/* ---- vectorAdd kernel variants (0 = the measured-best default) ------------------- */
#define GPUPROBE_VADD_AUTO 0
#define GPUPROBE_VADD_LDG128_U2 1  /* 128-bit ld/st, 2 vectors of each input in flight per thread */
#define GPUPROBE_VADD_LDG128_U4 2  /* 128-bit ld/st, 4 in flight                                   */
#define GPUPROBE_VADD_LDG128_U8 3  /* 128-bit ld/st, 8 in flight                                   */
#define GPUPROBE_VADD_LDG256_U2 4  /* 256-bit ld/st (sm_100 v8.f32), 2 in flight                   */
#define GPUPROBE_VADD_LDG256_U4 5  /* 256-bit ld/st, 4 in flight                                   */
#define GPUPROBE_VADD_TMA 6        /* cp.async.bulk ring (16 KiB tiles) through shared memory, one persistent CTA per SM */
#define GPUPROBE_VADD_PERSIST128 7 /* 128-bit ld/st, grid = k * SM count, grid-stride tiles        */
#define GPUPROBE_VADD_SCALAR 8     /* 32-bit ld/st; the unaligned / ragged-tail path                */
#define GPUPROBE_VADD_TMA_8K 9     /* as TMA with 8 KiB tiles, two CTAs per SM                     */
#define GPUPROBE_VADD_LDG256_U2_EF 10 /* 256-bit ld/st with L2::evict_first on loads and stores    */
#define GPUPROBE_VADD_LDG128_U1 11 /* 128-bit ld/st, 1 vector of each input per thread: 4 KiB tiles  */
#define GPUPROBE_VADD_LDG256_U1 12 /* 256-bit ld/st, 1 vector of each input per thread: 8 KiB tiles  */
#define GPUPROBE_VADD_SMALL 13 /* one launch, bounds-checked 128-bit ld/st: AUTO's choice for ragged n < 2^22 */
#define GPUPROBE_VADD_NUM_VARIANTS 14

typedef struct gpuprobe_devinfo {
    char name[128];
    int cc_major, cc_minor;
    int sm_count;
    int mem_clock_khz;
    int mem_bus_bits;
    int max_smem_optin; /* bytes of shared memory a CTA can opt in to */
    uint64_t hbm_bytes;
    uint64_t l2_bytes;
} gpuprobe_devinfo;

typedef struct gpuprobe_vadd_cfg {
    int dev;          /* CUDA ordinal; the sample always used device 0                          */
    uint64_t n_elems; /* fp32 elements per array; the sample used 50,000 (BASELINE.json cfg 0)  */
    int iters;        /* timed launches (>=1)                                                   */
    int warmup;       /* untimed launches before them                                           */
    uint32_t seed;
    int gen_kind;     /* GPUPROBE_GEN_*                                                         */
    int variant;      /* GPUPROBE_VADD_*                                                        */
    int verify;       /* bit0: digest of C; bit1: on-device regenerate-and-compare (sanity)     */
    double peak_nominal_gbps;  /* 0 -> 8000 (north_star's nominal B200 HBM figure)              */
    double peak_measured_gbps; /* 0 -> unknown, frac_measured reported as 0                     */
} gpuprobe_vadd_cfg;

typedef struct gpuprobe_vadd_result {
    uint64_t n;
    uint64_t sum64;      /* sum of the uint32 bit patterns of C, mod 2^64 */
    uint32_t xor32;      /* xor of the same                               */
    uint32_t variant;    /* variant that ran                              */
    uint64_t mismatches; /* elements where C != regenerated a+b (sanity)  */
    uint64_t algo_bytes; /* 12 * n                                        */
    double ms_best, ms_median; /* per launch, each launch inside its own CUDA-event pair          */
    double ms_mean;            /* per launch, `iters` launches back to back inside one event pair */
    double gbps_best, gbps_median;
    double frac_nominal, frac_measured; /* of gbps_median */
    int l2_resident;     /* 1 if 12*n fits in L2: not an HBM measurement  */
    int launches;        /* kernel launches per iteration (1, or 2 with a ragged tail) */
} gpuprobe_vadd_result;

#define GPUPROBE_P2P_SERIAL 0 /* one directed pair at a time                                  */
#define GPUPROBE_P2P_ROUNDS 1 /* ndev-1 rounds, dst = (src + r) mod ndev, all pairs of a round concurrently */
#define GPUPROBE_P2P_ROUNDS_PUSH 2 /* same schedule, but the kernel runs on src and STORES into dst over NVLink */
#define GPUPROBE_P2P_SERIAL_PUSH 3 /* one directed pair at a time, kernel on src storing into dst               */
/* The same schedules with the destination's COPY ENGINE moving the bytes (cudaMemcpyPeerAsync) and the checksum taken
 * from the landed copy by the digest kernel afterwards.  Not the probe north_star describes (no SM touches the link), but
 * the reference point a validator needs beside it: DMA transfers use larger link packets than SM loads/stores can, so
 * they show what the links themselves carry (about 770 GB/s per direction even with every GPU sending and receiving). */
#define GPUPROBE_P2P_SERIAL_CE 4
#define GPUPROBE_P2P_ROUNDS_CE 5

typedef struct gpuprobe_p2p_cfg {
    int ndev;
    const int *devs;         /* NULL -> 0..ndev-1 */
    uint64_t bytes_per_pair; /* multiple of 16     */
    int iters, warmup;
    int mode;                /* GPUPROBE_P2P_*     */
    uint32_t seed;
    double peak_gbps;        /* 0 -> 900 (NVLink 5 per direction) */
} gpuprobe_p2p_cfg;

typedef struct gpuprobe_p2p_cell {
    int src, dst;       /* CUDA ordinals; data flows src -> dst, the kernel runs on dst */
    int peer_ok;        /* cudaDeviceCanAccessPeer(dst, src)                            */
    int verified;       /* checksum_got == checksum_expected && copy intact             */
    uint64_t checksum_expected, checksum_got; /* sum of uint32 words mod 2^64           */
    uint32_t xor_expected, xor_got;
    double ms_best, ms_median;
    double gbps;        /* bytes_per_pair / ms_median                                   */
    double frac_of_peak;
} gpuprobe_p2p_cell;

typedef struct gpuprobe_ar_cfg {
    int ndev;        /* 2, 4 or 8 (any >= 2 accepted) */
    const int *devs; /* NULL -> 0..ndev-1              */
    uint64_t count;  /* int32 elements per rank        */
    int iters, warmup;
    uint32_t seed;
    double peak_gbps; /* 0 -> 900 */
} gpuprobe_ar_cfg;

typedef struct gpuprobe_ar_result {
    int ndev;
    int nccl_version;
    uint64_t bytes;      /* count * 4 per rank          */
    uint64_t mismatches; /* over all ranks; must be 0   */
    double ms_best, ms_median; /* max over ranks per iteration */
    double algbw_gbps;   /* bytes / ms_median           */
    double busbw_gbps;   /* algbw * 2(n-1)/n            */
    double frac_of_peak;
    /* ABI 2: what the ranks hold after the last in-place round, so that a caller (the parity tests, the
     * Go validator) can compare the collective's OUTPUT with a value computed elsewhere instead of
     * trusting `mismatches`, which the library derives from its own generator. */
    uint64_t sum64;      /* sum of rank 0's words mod 2^64                                        */
    uint32_t xor32;      /* xor of the same                                                       */
    uint32_t rounds;     /* in-place rounds run: every word is ndev^(rounds-1) * sum_r x_r[i]     */
    int ranks_equal;     /* 1 if every rank's buffer has the digest of rank 0's                   */
    int impl;            /* GPUPROBE_AR_IMPL_* that produced the result                           */
} gpuprobe_ar_result;

#define GPUPROBE_AR_IMPL_NCCL 0
#define GPUPROBE_AR_IMPL_NVLS 1
#define GPUPROBE_AR_IMPL_P2P 2    /* two-shot: pull-reduce the slice, push the sum            */
#define GPUPROBE_AR_IMPL_HYBRID 3
#define GPUPROBE_AR_IMPL_PUSH 4   /* two-shot with posted writes only (scatter, add, broadcast) */
#define GPUPROBE_AR_IMPL_CE 5     /* two-shot with the copy engines moving the data, a kernel adding */
#define GPUPROBE_AR_IMPL_NVLS_CE 6 /* the multimem kernel on the head of the buffer, the copy-engine two-shot on the tail */

/* ---- library / device queries --------------------------------------------------- */

/* No reference counterpart (the sample binary has no version handshake). */
GPUPROBE_API int gpuprobe_abi_version(void);

/* Stands in for the sample's implicit cudaGetDeviceCount via cudaMalloc on device 0; the
 * cuda workload pod sees all GPUs (validator/cuda-workload-validation.yaml:21-25), the
 * plugin pod the one it was allocated (validator/main.go:973-983). */
GPUPROBE_API int gpuprobe_device_count(int *n);
GPUPROBE_API int gpuprobe_device_info(int dev, gpuprobe_devinfo *out);
GPUPROBE_API const char *gpuprobe_strerror(int code);

/* ---- (a) vectorAdd ---------------------------------------------------------------- */

/* The whole probe: allocate, generate A and B on the device, run C = A + B `warmup+iters`
 * times, digest and sanity-check C, free.  Replaces the body of the external `vectorAdd`
 * binary (installed by validator/Dockerfile:33-35,49; run at
 * validator/cuda-workload-validation.yaml:19-20). */
GPUPROBE_API int gpuprobe_vectoradd(const gpuprobe_vadd_cfg *cfg, gpuprobe_vadd_result *out, char *err, size_t errlen);

/* Caller-supplied HOST inputs: copies a,b to the device (chunked, pipelined over separate
 * copy streams), adds, copies C back, so the CALLER's host loop does the bit-exact compare
 * -- the sample's h_A/h_B -> cudaMemcpy -> kernel -> cudaMemcpy -> h_C sequence.  Pinned,
 * device-mapped buffers (gpuprobe_host_alloc) are not staged at all: the add kernel loads A
 * and B from host memory over PCIe and stores C back over PCIe (env GPUPROBE_BUF_ZEROCOPY=0
 * forces the staged pipeline); pageable ones go through the library's pinned bounce buffers
 * and the chunked H2D / kernel / D2H pipeline. */
GPUPROBE_API int gpuprobe_vectoradd_buffers(int dev, const float *a, const float *b, float *c, uint64_t n, char *err,
                               size_t errlen);

/* Caller-supplied DEVICE buffers: enqueues the add kernel(s) on `stream` (a cudaStream_t,
 * NULL = legacy default stream) and returns without synchronising.  Allocates nothing.
 * This is what bench.py times with CUDA events. */
GPUPROBE_API int gpuprobe_vectoradd_device(int dev, const float *d_a, const float *d_b, float *d_c, uint64_t n, int variant,
                              void *stream, char *err, size_t errlen);

/* The deterministic generator on the host, exported so that Go / C++ / Python callers all
 * produce the inputs the device generator produces (stands in for the sample's
 * rand()/(float)RAND_MAX fill of h_A, h_B).  a[j], b[j] receive elements first+j. */
GPUPROBE_API void gpuprobe_fill_inputs(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int gen_kind);

/* Same generator on the device, into caller-supplied device buffers (either may be NULL);
 * asynchronous on `stream`. */
GPUPROBE_API int gpuprobe_fill_inputs_device(int dev, float *d_a, float *d_b, uint64_t first, uint64_t n, uint32_t seed,
                                int gen_kind, void *stream, char *err, size_t errlen);

/* sum64 / xor32 of the bit patterns of a device array of n 32-bit words; synchronises
 * `stream` before returning.  Uses the device's cached scratch buffer (allocated at the
 * first digest or verify on that device, freed by gpuprobe_shutdown). */
GPUPROBE_API int gpuprobe_digest_device(int dev, const void *d_words, uint64_t n, uint64_t *sum64, uint32_t *xor32,
                           void *stream, char *err, size_t errlen);

/* The same digest for callers that must not allocate: a re-check that runs while tenant jobs
 * hold (nearly) all of HBM or the whole MIG slice, and any caller that may not disturb those
 * jobs -- cudaMalloc/cudaFree take device-wide locks and cudaFree waits for the device to
 * idle -- or that lives inside a framework's memory pool and may not go around it.
 * No counterpart in the reference (which never inspects device data at all).
 * Scratch is a fixed GPUPROBE_NOALLOC_SCRATCH_BYTES array inside the library image, resident on
 * every GPU that loads the library until the process exits.  It allows at most
 * GPUPROBE_NOALLOC_MAX_WORDS words per call; more is rejected with GPUPROBE_EINVAL.  The call
 * performs no cudaMalloc/cudaFree, clears the 16 bytes of state that must start at zero on
 * `stream`, launches, copies the 16-byte result back and synchronises `stream`.  All calls on
 * a device share the array: the library serialises them, and work that reads the array must
 * not overlap on the GPU.  The first call on each device looks the array's address up (set-up);
 * make that first call outside any stream capture.  With CUDA's default lazy module loading that
 * look-up may also be what loads the library's device code onto the GPU, and the array is reserved
 * at that moment -- so "allocates nothing" holds from the second call on: make the set-up call at
 * start-up (n = 0 is enough), before the GPU fills up, not when memory is already exhausted. */
#define GPUPROBE_NOALLOC_MAX_WORDS (1ull << 30)
#define GPUPROBE_NOALLOC_SCRATCH_BYTES 196864
GPUPROBE_API int gpuprobe_digest_device_noalloc(int dev, const void *d_words, uint64_t n, uint64_t *sum64,
                                                uint32_t *xor32, void *stream, char *err, size_t errlen);

/* Counts elements of d_c whose bits differ from regenerated a[first+j] + b[first+j];
 * synchronises `stream`.  Same adder as the probe, so a sanity check, not parity. */
GPUPROBE_API int gpuprobe_verify_device(int dev, const float *d_c, uint64_t first, uint64_t n, uint32_t seed, int gen_kind,
                           uint64_t *mismatches, void *stream, char *err, size_t errlen);

/* Pinned host memory for gpuprobe_vectoradd_buffers callers (cudaHostAlloc / cudaFreeHost). */
GPUPROBE_API int gpuprobe_host_alloc(void **p, size_t bytes);
GPUPROBE_API int gpuprobe_host_free(void *p);

/* ---- (b) NVLink P2P all-pairs copy + checksum --------------------------------------- */

/* New capability (the reference never touches a second GPU).  cells is caller-allocated,
 * ndev*ndev entries, row-major [src][dst]; the diagonal is left zeroed.  Returns
 * GPUPROBE_FAIL_VERIFY if any pair's checksum is wrong, GPUPROBE_ENOPEER if a pair has
 * no peer access (its cell has peer_ok = 0). */
GPUPROBE_API int gpuprobe_p2p_matrix(const gpuprobe_p2p_cfg *cfg, gpuprobe_p2p_cell *cells, char *err, size_t errlen);

/* The same two kernels (kernel: 0 = cp.async.bulk ring, 1 = register path) with source and landing buffer on ONE device.
 * No link is crossed, so this is not a fabric figure (cell->gbps is local HBM copy bandwidth, frac_of_peak stays 0); it
 * lets the kernels and their checksum reduction be checked against the expected value where only one GPU is visible --
 * the plugin workload's pod (validator/main.go:973-983 allocates it one GPU) or a single-GPU test box.  New capability. */
GPUPROBE_API int gpuprobe_p2p_loopback(int dev, uint64_t bytes, int iters, int kernel, uint32_t seed, gpuprobe_p2p_cell *cell,
                                       char *err, size_t errlen);

/* Expected checksum of source `src`'s buffer, computed on the host (threads <= 0: all). */
GPUPROBE_API void gpuprobe_p2p_expected(uint64_t nwords, uint32_t seed, int src, int threads, uint64_t *sum64,
                           uint32_t *xor32);

/* ---- (c) NCCL int32 allreduce fabric check ------------------------------------------ */

/* New capability.  Single process, one communicator over cfg->ndev devices
 * (ncclCommInitAll); NCCL is loaded at run time (env GPUPROBE_NCCL_LIB, else libnccl.so.2). */
GPUPROBE_API int gpuprobe_allreduce_i32(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c') The same check as one hand-written kernel per GPU over NVLink multicast (NVLS): fill,
 * multimem.ld_reduce (the switch adds) + multimem.st (the switch broadcasts) and verify in a
 * single cooperative launch, barriers through a multicast flag word.  Same cfg/result as the
 * NCCL probe (nccl_version is 0).  Returns GPUPROBE_ENOPEER if a device lacks multicast support.
 * New capability, no reference counterpart. */
GPUPROBE_API int gpuprobe_allreduce_i32_nvls(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c'') Same launch structure and buffers as the NVLS kernel, but the data moves over plain peer
 * mappings: each rank loads its 1/n slice from every rank (n-1 loads cross NVLink), adds, and
 * stores the sum into every rank's buffer (n-1 stores cross NVLink).  ndev must be 2, 4 or 8.
 * The better choice at 2 ranks, where the switch path moves 1.5x the bytes; no reference counterpart.
 * Works without NVLS: when a device lacks multicast support (or env GPUPROBE_NO_MULTICAST=1) the
 * cross-GPU barriers go through per-rank flag slots written with peer stores instead of multimem.red. */
GPUPROBE_API int gpuprobe_allreduce_i32_p2p(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c''') Hybrid of the two: the head of each rank's slice goes through the switch (multimem), the
 * tail over peer mappings, on disjoint CTAs of the same kernel, so the switch's reduction engine and
 * the link bandwidth it leaves idle are used together.  ndev must be 2, 4 or 8.  The split is set by
 * env GPUPROBE_HYBRID_PERMILLE (share of the data through the switch, default 950) and
 * GPUPROBE_HYBRID_BLOCK_PERMILLE (share of the CTAs on that part, default 350). */
GPUPROBE_API int gpuprobe_allreduce_i32_hybrid(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c'''') All-push two-shot: every rank first stores its copy of each slice into a staging area in the slice
 * owner's memory (peer stores), the owner adds the nd copies out of its own HBM and stores the sum into every
 * rank's buffer (peer stores again).  No load crosses NVLink, so no read-request packets compete with the data on
 * the opposite direction of the links; costs (nd-1)/nd of the payload in extra device memory per rank and a
 * third in-kernel synchronisation.  ndev must be 2, 4 or 8; works with or without multicast.  No reference counterpart. */
GPUPROBE_API int gpuprobe_allreduce_i32_push(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c-ce) Two-shot with every transfer done by the GPUs' copy engines (cudaMemcpyPeerAsync, ordered by events across
 * the devices' streams) and only the addition in a kernel, chunked so that the add overlaps the transfers.  Exists because
 * of a measurement (profiles/r02_fabric.md): SM-issued loads and stores cross NVLink in 128-byte packets and top out at
 * 667-700 GB/s of payload per direction when both directions are busy, DMA transfers reach 770.  Any ndev >= 2; needs
 * peer access, not multicast; uses (ndev-1)/ndev of the payload in staging memory per rank.  No reference counterpart. */
GPUPROBE_API int gpuprobe_allreduce_i32_ce(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* (c-nvls-ce) The switch and the copy engines at once: the multimem kernel (c') reduces the first
 * GPUPROBE_NVLSCE_PERMILLE/1000 of the buffer (default 800) while the copy-engine two-shot (c-ce) reduces the rest, each
 * on its own streams of every GPU.  The switch's reduction path sustains about 500 GB/s of link traffic per direction and
 * leaves link time idle; DMA packets fill it better than SM-issued ones.  Needs multicast; any ndev >= 2.
 * No reference counterpart. */
GPUPROBE_API int gpuprobe_allreduce_i32_nvls_ce(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* What the validator runs when GPUPROBE_AR_IMPL does not name an implementation: the measured best for the rank count
 * and size (profiles/r02_fabric.md) -- the copy-engine two-shot for large buffers at 2 ranks (>= 32 MiB) and at 3-4 ranks
 * (>= 512 MiB), otherwise the all-push kernel at 2 ranks, the hybrid kernel at 4 and 8, NVLS at other counts; the
 * peer-mapping two-shot or the copy engines where NVLink multicast is unavailable.  out->impl says which ran.
 * No reference counterpart. */
GPUPROBE_API int gpuprobe_allreduce_i32_auto(const gpuprobe_ar_cfg *cfg, gpuprobe_ar_result *out, char *err, size_t errlen);

/* Releases cached per-device state: streams, the digest/verify scratch, the staging slots and pinned
 * bounce buffers of gpuprobe_vectoradd_buffers.  Optional; the library re-creates them on demand. */
GPUPROBE_API void gpuprobe_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* GPUPROBE_H */

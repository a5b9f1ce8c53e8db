/*
 * gpuprobe_oracle.c -- CPU oracle (plain C). See gpuprobe_oracle.h for the
 * provenance statement: TEST INFRASTRUCTURE ONLY, PARITY UNPINNED.
 *
 * Build: make -C oracle   (gcc -O2, no -ffast-math, no -march flags that
 * would change rounding; SSE2 scalar/vector fp32 add is IEEE single rounding).
 */
#include "gpuprobe_oracle.h"

#include <math.h>
#include <pthread.h>
#include <string.h>
#include <xmmintrin.h>

/* ------------------------------------------------------------------ hash */

uint32_t oracle_mix32(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

/* DESIGN.md "Generator": h0 = mix32(seed ^ (stream+1)*0x9E3779B9);
 * h1 = mix32(h0 ^ hi32(i)); h = mix32(h1 ^ lo32(i)). */
uint32_t oracle_hash(uint32_t seed, uint32_t stream, uint64_t i)
{
    uint32_t h = oracle_mix32(seed ^ ((stream + 1u) * 0x9E3779B9U));
    h = oracle_mix32(h ^ (uint32_t)(i >> 32));
    h = oracle_mix32(h ^ (uint32_t)(i & 0xffffffffU));
    return h;
}

/* ------------------------------------------------------------- generator */

static const uint32_t EDGE_A[16] = {
    0x00000000U, 0x80000000U, 0x00000001U, 0x80000001U, /* +-0, +-min denormal        */
    0x007fffffU, 0x807fffffU, 0x00800000U, 0x80800000U, /* +-max denormal, +-min normal */
    0x3f7fffffU, 0x3f800001U, 0x3f800000U, 0xbf800000U, /* 1-2^-24, 1+2^-23, 1, -1     */
    0x7f7fffffU, 0xff7fffffU, 0x7f800000U, 0xff800000U, /* +-max finite, +-inf         */
};
/* B holds finite values only, so A+B is never inf + -inf (NaN payloads differ
 * legitimately between x86 and the GPU: SURVEY.md section 7 "Bit-exactness caveats"). */
static const uint32_t EDGE_B[16] = {
    0x00000000U, 0x80000000U, 0x00000001U, 0x80000001U,
    0x007fffffU, 0x807fffffU, 0x00800000U, 0x80800000U,
    0x33800000U, 0x33000000U, 0xb3800000U, 0xbf7fffffU, /* 2^-24, 2^-25, -2^-24, -(1-2^-24) */
    0xbf800001U, 0x7f7fffffU, 0xff7fffffU, 0x73000000U, /* -(1+2^-23), +-max finite, 2^103  */
};

uint32_t oracle_gen_bits(uint32_t seed, uint32_t stream, uint64_t i, int kind)
{
    uint32_t h = oracle_hash(seed, stream, i);
    switch (kind) {
    case ORACLE_GEN_EDGE:
        return (stream == ORACLE_STREAM_A ? EDGE_A : EDGE_B)[h & 15u];
    case ORACLE_GEN_BITS:
        if (((h >> 23) & 0xffu) == 0xffu)
            h ^= 0x00800000U; /* exponent 0xff -> 0xfe: keep it finite */
        return h;
    default: { /* ORACLE_GEN_UNIT */
        float f = (float)(h >> 8) * (1.0f / 16777216.0f); /* exact: 24-bit integer * 2^-24 */
        uint32_t u;
        memcpy(&u, &f, 4);
        return u;
    }
    }
}

void oracle_fill_inputs(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int kind)
{
    for (uint64_t j = 0; j < n; ++j) {
        uint32_t ua = oracle_gen_bits(seed, ORACLE_STREAM_A, first + j, kind);
        uint32_t ub = oracle_gen_bits(seed, ORACLE_STREAM_B, first + j, kind);
        memcpy(&a[j], &ua, 4);
        memcpy(&b[j], &ub, 4);
    }
}

struct fill_job {
    float *a, *b;
    uint64_t first, n;
    uint32_t seed;
    int kind;
};

static void *fill_worker(void *p)
{
    struct fill_job *j = (struct fill_job *)p;
    oracle_fill_inputs(j->a, j->b, j->first, j->n, j->seed, j->kind);
    return NULL;
}

void oracle_fill_inputs_mt(float *a, float *b, uint64_t first, uint64_t n, uint32_t seed, int kind, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if (threads == 1 || n < 65536u) {
        oracle_fill_inputs(a, b, first, n, seed, kind);
        return;
    }
    pthread_t tid[256];
    struct fill_job job[256];
    uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    int started = 0;
    for (int t = 0; t < threads; ++t) {
        uint64_t lo = per * (uint64_t)t;
        if (lo >= n) break;
        uint64_t hi = lo + per < n ? lo + per : n;
        job[t].a = a + lo;
        job[t].b = b + lo;
        job[t].first = first + lo;
        job[t].n = hi - lo;
        job[t].seed = seed;
        job[t].kind = kind;
        pthread_create(&tid[t], NULL, fill_worker, &job[t]);
        ++started;
    }
    for (int t = 0; t < started; ++t)
        pthread_join(tid[t], NULL);
}

/* --------------------------------------------------------- the host loop */

void oracle_vectoradd(const float *a, const float *b, float *c, uint64_t n)
{
    for (uint64_t i = 0; i < n; ++i)
        c[i] = a[i] + b[i];
}

struct vadd_job {
    const float *a, *b;
    float *c;
    uint64_t n;
};

static void *vadd_worker(void *p)
{
    struct vadd_job *j = (struct vadd_job *)p;
    oracle_vectoradd(j->a, j->b, j->c, j->n);
    return NULL;
}

#define ORACLE_MAX_THREADS 256

void oracle_vectoradd_mt(const float *a, const float *b, float *c, uint64_t n, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > ORACLE_MAX_THREADS) threads = ORACLE_MAX_THREADS;
    if (threads == 1 || n < (uint64_t)threads * 4096u) {
        oracle_vectoradd(a, b, c, n);
        return;
    }
    pthread_t tid[ORACLE_MAX_THREADS];
    struct vadd_job job[ORACLE_MAX_THREADS];
    uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    per = (per + 15u) & ~(uint64_t)15u;
    int started = 0;
    for (int t = 0; t < threads; ++t) {
        uint64_t lo = per * (uint64_t)t;
        if (lo >= n) break;
        uint64_t hi = lo + per < n ? lo + per : n;
        job[t].a = a + lo;
        job[t].b = b + lo;
        job[t].c = c + lo;
        job[t].n = hi - lo;
        pthread_create(&tid[t], NULL, vadd_worker, &job[t]);
        ++started;
    }
    for (int t = 0; t < started; ++t)
        pthread_join(tid[t], NULL);
}

/* --------------------------------------------------------------- digests */

void oracle_digest(const float *c, uint64_t n, uint64_t *sum64, uint32_t *xor32)
{
    uint64_t s = 0;
    uint32_t x = 0;
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t u;
        memcpy(&u, &c[i], 4);
        s += u;
        x ^= u;
    }
    *sum64 = s;
    *xor32 = x;
}

struct range_job {
    uint64_t first, n;
    uint32_t seed;
    int kind, src;
    uint64_t sum;
    uint32_t xr;
};

static void *vadd_digest_worker(void *p)
{
    struct range_job *j = (struct range_job *)p;
    uint64_t s = 0;
    uint32_t x = 0;
    for (uint64_t i = j->first; i < j->first + j->n; ++i) {
        uint32_t ua = oracle_gen_bits(j->seed, ORACLE_STREAM_A, i, j->kind);
        uint32_t ub = oracle_gen_bits(j->seed, ORACLE_STREAM_B, i, j->kind);
        float fa, fb, fc;
        memcpy(&fa, &ua, 4);
        memcpy(&fb, &ub, 4);
        fc = fa + fb;
        uint32_t uc;
        memcpy(&uc, &fc, 4);
        s += uc;
        x ^= uc;
    }
    j->sum = s;
    j->xr = x;
    return NULL;
}

static void *p2p_digest_worker(void *p)
{
    struct range_job *j = (struct range_job *)p;
    uint64_t s = 0;
    uint32_t x = 0;
    for (uint64_t i = j->first; i < j->first + j->n; ++i) {
        uint32_t w = oracle_hash(j->seed, ORACLE_STREAM_P2P(j->src), i);
        s += w;
        x ^= w;
    }
    j->sum = s;
    j->xr = x;
    return NULL;
}

static void run_range(void *(*fn)(void *), uint64_t first, uint64_t n, uint32_t seed, int kind,
                      int src, int threads, uint64_t *sum64, uint32_t *xor32)
{
    if (threads < 1) threads = 1;
    if (threads > ORACLE_MAX_THREADS) threads = ORACLE_MAX_THREADS;
    if (n < 65536u) threads = 1;
    pthread_t tid[ORACLE_MAX_THREADS];
    struct range_job job[ORACLE_MAX_THREADS];
    uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    int started = 0;
    for (int t = 0; t < threads; ++t) {
        uint64_t lo = per * (uint64_t)t;
        if (lo >= n && !(n == 0 && t == 0)) break;
        uint64_t hi = lo + per < n ? lo + per : n;
        job[t].first = first + lo;
        job[t].n = hi - lo;
        job[t].seed = seed;
        job[t].kind = kind;
        job[t].src = src;
        job[t].sum = 0;
        job[t].xr = 0;
        if (threads == 1)
            fn(&job[t]);
        else
            pthread_create(&tid[t], NULL, fn, &job[t]);
        ++started;
    }
    uint64_t s = 0;
    uint32_t x = 0;
    for (int t = 0; t < started; ++t) {
        if (threads != 1) pthread_join(tid[t], NULL);
        s += job[t].sum;
        x ^= job[t].xr;
    }
    *sum64 = s;
    *xor32 = x;
}

void oracle_vectoradd_digest(uint64_t first, uint64_t n, uint32_t seed, int kind, int threads,
                             uint64_t *sum64, uint32_t *xor32)
{
    run_range(vadd_digest_worker, first, n, seed, kind, 0, threads, sum64, xor32);
}

/* ------------------------------------------------ the sample's own check */

uint64_t oracle_sample_check(const float *a, const float *b, const float *c, uint64_t n)
{
    /* published cuda-samples vectorAdd: if (fabs(h_A[i] + h_B[i] - h_C[i]) > 1e-5) fail */
    uint64_t bad = 0;
    for (uint64_t i = 0; i < n; ++i) {
        double d = fabs((double)(float)(a[i] + b[i] - c[i]));
        if (d > 1e-5) ++bad;
    }
    return bad;
}

uint64_t oracle_count_mismatch(const float *x, const float *y, uint64_t n)
{
    uint64_t bad = 0;
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t ux, uy;
        memcpy(&ux, &x[i], 4);
        memcpy(&uy, &y[i], 4);
        bad += (ux != uy);
    }
    return bad;
}

/* ---------------------------------------------------------- fabric probes */

void oracle_p2p_fill(uint32_t *w, uint64_t first, uint64_t n, uint32_t seed, int src)
{
    for (uint64_t j = 0; j < n; ++j)
        w[j] = oracle_hash(seed, ORACLE_STREAM_P2P(src), first + j);
}

void oracle_p2p_checksum(uint64_t nwords, uint32_t seed, int src, int threads, uint64_t *sum64,
                         uint32_t *xor32)
{
    run_range(p2p_digest_worker, 0, nwords, seed, 0, src, threads, sum64, xor32);
}

void oracle_allreduce_fill(int32_t *x, uint64_t first, uint64_t n, uint32_t seed, int rank)
{
    for (uint64_t j = 0; j < n; ++j)
        x[j] = (int32_t)oracle_hash(seed, ORACLE_STREAM_AR(rank), first + j);
}

/* Digest of what every rank must hold after `rounds` in-place int32 sum-allreduces over `nranks` ranks:
 * word i = nranks^(rounds-1) * sum_r x_r[i] in wraparound 32-bit arithmetic (each further in-place round adds
 * nranks equal copies). Threaded because the probe's full size is 2^28 words x 8 ranks of hashing. */
static void *allreduce_digest_worker(void *p)
{
    struct range_job *j = (struct range_job *)p;
    uint32_t mult = 1;
    for (int k = 1; k < j->kind; ++k) /* kind carries `rounds` */
        mult *= (uint32_t)j->src;     /* src carries `nranks` */
    uint64_t s = 0;
    uint32_t x = 0;
    for (uint64_t i = j->first; i < j->first + j->n; ++i) {
        uint32_t w = 0;
        for (int r = 0; r < j->src; ++r)
            w += oracle_hash(j->seed, ORACLE_STREAM_AR(r), i);
        w *= mult;
        s += w;
        x ^= w;
    }
    j->sum = s;
    j->xr = x;
    return NULL;
}

void oracle_allreduce_digest(uint64_t first, uint64_t n, uint32_t seed, int nranks, int rounds, int threads,
                             uint64_t *sum64, uint32_t *xor32)
{
    run_range(allreduce_digest_worker, first, n, seed, rounds < 1 ? 1 : rounds, nranks, threads, sum64, xor32);
}

void oracle_allreduce_expected(int32_t *out, uint64_t first, uint64_t n, uint32_t seed, int nranks)
{
    for (uint64_t j = 0; j < n; ++j) {
        uint32_t s = 0; /* unsigned wraparound == two's-complement int32 sum */
        for (int r = 0; r < nranks; ++r)
            s += oracle_hash(seed, ORACLE_STREAM_AR(r), first + j);
        out[j] = (int32_t)s;
    }
}

/* ---------------------------------------------------------------- fp env */

int oracle_fp_env_ok(void)
{
    unsigned int csr = _mm_getcsr();
    /* bit 15 FTZ, bit 6 DAZ, bits 13-14 rounding control (00 = nearest) */
    return (csr & 0x8040u) == 0 && (csr & 0x6000u) == 0;
}

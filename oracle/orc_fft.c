/* orc_fft.c -- negacyclic f64 FFT wrapper of the reference (TEST INFRASTRUCTURE, see tfhe_oracle.h).
 *
 * Restates fft_impl/fft64/math/fft/mod.rs: fold N reals into M = N/2 complex, multiply by the
 * twist e^{i pi j / N} (Twisties::new :58-69), complex DFT of size M (concrete-fft 0.3.0
 * `unordered::Plan`, un-vendored: its output ordering is unspecified and only has to agree
 * between fwd and inv -- here: decimation-in-frequency forward leaves bit-reversed order,
 * decimation-in-time inverse consumes it).  Pointwise products are order-agnostic. */
#include "tfhe_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

typedef struct { int N; double *twist; double *w; } plan_t; /* twist[2M], w[2*(M/2)] = exp(-2 pi i t / M) */
static plan_t g_plans[32];
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static const plan_t *get_plan(int N) {
    int lg = 0; while ((1 << lg) < N) lg++;
    plan_t *p = &g_plans[lg];
    if (__atomic_load_n(&p->N, __ATOMIC_ACQUIRE) == N) return p;
    pthread_mutex_lock(&g_lock);
    if (p->N != N) {
        int M = N / 2;
        double *tw = (double *)malloc(sizeof(double) * 2 * M);
        double *w = (double *)malloc(sizeof(double) * 2 * (M / 2 > 0 ? M / 2 : 1));
        const long double PI = 3.14159265358979323846264338327950288L;
        for (int j = 0; j < M; j++) {
            long double a = PI * (long double)j / (long double)N;
            tw[2 * j] = (double)cosl(a); tw[2 * j + 1] = (double)sinl(a);
        }
        for (int t = 0; t < M / 2; t++) {
            long double a = -2.0L * PI * (long double)t / (long double)M;
            w[2 * t] = (double)cosl(a); w[2 * t + 1] = (double)sinl(a);
        }
        p->twist = tw; p->w = w;
        __atomic_store_n(&p->N, N, __ATOMIC_RELEASE);
    }
    pthread_mutex_unlock(&g_lock);
    return p;
}

/* in-place DIF radix-2, natural in -> bit-reversed out, kernel exp(-2 pi i jk/M) */
static void fft_dif(double *x, int M, const double *w) {
    for (int len = M; len >= 2; len >>= 1) {
        int half = len >> 1, step = M / len;
        for (int b = 0; b < M; b += len) {
            for (int q = 0; q < half; q++) {
                double *u = x + 2 * (b + q), *v = x + 2 * (b + q + half);
                double ar = u[0], ai = u[1], br = v[0], bi = v[1];
                double wr = w[2 * q * step], wi = w[2 * q * step + 1];
                double dr = ar - br, di = ai - bi;
                u[0] = ar + br; u[1] = ai + bi;
                v[0] = dr * wr - di * wi; v[1] = dr * wi + di * wr;
            }
        }
    }
}
/* in-place DIT radix-2 inverse (unnormalised), bit-reversed in -> natural out, kernel exp(+2 pi i jk/M) */
static void fft_dit_inv(double *x, int M, const double *w) {
    for (int len = 2; len <= M; len <<= 1) {
        int half = len >> 1, step = M / len;
        for (int b = 0; b < M; b += len) {
            for (int q = 0; q < half; q++) {
                double *u = x + 2 * (b + q), *v = x + 2 * (b + q + half);
                double wr = w[2 * q * step], wi = -w[2 * q * step + 1];
                double br = v[0] * wr - v[1] * wi, bi = v[0] * wi + v[1] * wr;
                double ar = u[0], ai = u[1];
                u[0] = ar + br; u[1] = ai + bi;
                v[0] = ar - br; v[1] = ai - bi;
            }
        }
    }
}

static inline double to_signed_f64(uint64_t v, int bits) {
    /* into_signed().cast_into() : fft/mod.rs:228-229 */
    if (bits == 64) return (double)(int64_t)v;
    return (double)(int32_t)(uint32_t)v;
}

/* fft/mod.rs:220-239 (convert_forward_integer) + :496-515 */
void orc_fft_forward_integer(int N, const uint64_t *poly, int bits, double *out) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    for (int j = 0; j < M; j++) {
        double re = to_signed_f64(poly[j], bits), im = to_signed_f64(poly[j + M], bits);
        double wr = p->twist[2 * j], wi = p->twist[2 * j + 1];
        out[2 * j] = re * wr - im * wi;
        out[2 * j + 1] = re * wi + im * wr;
    }
    fft_dif(out, M, p->w);
}

/* fft/mod.rs:197-218 (convert_forward_torus: normalisation 2^-BITS) */
void orc_fft_forward_torus(int N, const uint64_t *poly, int bits, double *out) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    double norm = ldexp(1.0, -bits);
    for (int j = 0; j < M; j++) {
        double re = to_signed_f64(poly[j], bits) * norm, im = to_signed_f64(poly[j + M], bits) * norm;
        double wr = p->twist[2 * j], wi = p->twist[2 * j + 1];
        out[2 * j] = re * wr - im * wi;
        out[2 * j + 1] = re * wi + im * wr;
    }
    fft_dif(out, M, p->w);
}

/* commons/math/torus/mod.rs:72-78 from_torus, followed by the signed cast of fft/mod.rs:285-304 */
static inline uint64_t from_torus(double t, int bits) {
    double fr = t - round(t);
    fr *= ldexp(1.0, bits);
    fr = round(fr);
    /* Rust `as i64` saturates; only reachable at the measure-zero tie fr == 2^63 */
    if (fr >= 9223372036854775807.0) return (uint64_t)INT64_MAX;
    if (fr <= -9223372036854775808.0) return (uint64_t)INT64_MIN;
    return (uint64_t)(int64_t)fr;
}

/* fft/mod.rs:285-326,539-557 : inverse FFT, * conj(twist) / M, round to torus, wrapping add */
void orc_fft_add_backward_torus(int N, uint64_t *poly, int bits, double *f) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    uint64_t m = bits == 64 ? ~0ULL : ((1ULL << bits) - 1);
    fft_dit_inv(f, M, p->w);
    double inv = 1.0 / (double)M;
    for (int j = 0; j < M; j++) {
        double wr = p->twist[2 * j], wi = -p->twist[2 * j + 1];
        double re = (f[2 * j] * wr - f[2 * j + 1] * wi) * inv;
        double im = (f[2 * j] * wi + f[2 * j + 1] * wr) * inv;
        poly[j] = (poly[j] + from_torus(re, bits)) & m;
        poly[j + M] = (poly[j + M] + from_torus(im, bits)) & m;
    }
}

/* fft/tests.rs:83-100 convolution_naive (negacyclic schoolbook, wrapping) */
void orc_negacyclic_schoolbook(int N, const uint64_t *a, const uint64_t *b, uint64_t *out, int bits) {
    uint64_t m = bits == 64 ? ~0ULL : ((1ULL << bits) - 1);
    memset(out, 0, sizeof(uint64_t) * N);
    for (int i = 0; i < N; i++) {
        uint64_t ai = a[i];
        for (int j = 0; j < N - i; j++) out[i + j] += ai * b[j];
        for (int j = N - i; j < N; j++) out[i + j - N] -= ai * b[j];
    }
    for (int i = 0; i < N; i++) out[i] &= m;
}

/* Exact a * s mod (X^N+1, 2^bits) for binary s: split a into 16-bit limbs, every limb product has
 * coefficients < 2^16 * N <= 2^30, far inside f64 exactness after rounding. Oracle-only helper for
 * GLWE encryption (algorithms/glwe_encryption.rs: body += sum_j A_j * S_j). */
void orc_negacyclic_mul_binary(int N, const uint64_t *a, const uint64_t *s, uint64_t *out, int bits) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    uint64_t m = bits == 64 ? ~0ULL : ((1ULL << bits) - 1);
    double *fs = (double *)malloc(sizeof(double) * 2 * M), *fa = (double *)malloc(sizeof(double) * 2 * M);
    uint64_t *limb = (uint64_t *)malloc(sizeof(uint64_t) * N);
    orc_fft_forward_integer(N, s, 64, fs);
    memset(out, 0, sizeof(uint64_t) * N);
    for (int sh = 0; sh < bits; sh += 16) {
        for (int j = 0; j < N; j++) limb[j] = (a[j] >> sh) & 0xFFFF;
        orc_fft_forward_integer(N, limb, 64, fa);
        for (int j = 0; j < M; j++) {
            double r = fa[2 * j] * fs[2 * j] - fa[2 * j + 1] * fs[2 * j + 1];
            double i = fa[2 * j] * fs[2 * j + 1] + fa[2 * j + 1] * fs[2 * j];
            fa[2 * j] = r; fa[2 * j + 1] = i;
        }
        fft_dit_inv(fa, M, p->w);
        double inv = 1.0 / (double)M;
        for (int j = 0; j < M; j++) {
            double wr = p->twist[2 * j], wi = -p->twist[2 * j + 1];
            double re = (fa[2 * j] * wr - fa[2 * j + 1] * wi) * inv;
            double im = (fa[2 * j] * wi + fa[2 * j + 1] * wr) * inv;
            out[j] += (uint64_t)(int64_t)llround(re) << sh;
            out[j + M] += (uint64_t)(int64_t)llround(im) << sh;
        }
    }
    for (int j = 0; j < N; j++) out[j] &= m;
    free(fs); free(fa); free(limb);
}

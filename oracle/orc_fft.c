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

/* Fourier buffers are structure-of-arrays: M real parts followed by M imaginary parts (N doubles per polynomial), so
 * that gcc vectorises the butterfly and pointwise loops (the reference gets its speed from pulp/AVX-512 intrinsics). */
typedef struct { int N; double *twr, *twi;   /* twist e^{i pi j/N}, j < M */
                 double *wr, *wi;            /* per-stage twiddles, stage `len` at offset len/2: exp(-2 pi i q/len), q < len/2 */
               } plan_t;
static plan_t g_plans[32];
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static const plan_t *get_plan(int N) {
    int lg = 0; while ((1 << lg) < N) lg++;
    plan_t *p = &g_plans[lg];
    if (__atomic_load_n(&p->N, __ATOMIC_ACQUIRE) == N) return p;
    pthread_mutex_lock(&g_lock);
    if (p->N != N) {
        int M = N / 2;
        double *twr = (double *)aligned_alloc(64, sizeof(double) * (M + 8)), *twi = (double *)aligned_alloc(64, sizeof(double) * (M + 8));
        double *wr = (double *)aligned_alloc(64, sizeof(double) * (M + 8)), *wi = (double *)aligned_alloc(64, sizeof(double) * (M + 8));
        const long double PI = 3.14159265358979323846264338327950288L;
        for (int j = 0; j < M; j++) {
            long double a = PI * (long double)j / (long double)N;
            twr[j] = (double)cosl(a); twi[j] = (double)sinl(a);
        }
        wr[0] = 1.0; wi[0] = 0.0;
        for (int len = 2; len <= M; len <<= 1)
            for (int q = 0; q < len / 2; q++) {
                long double a = -2.0L * PI * (long double)q / (long double)len;
                wr[len / 2 + q] = (double)cosl(a); wi[len / 2 + q] = (double)sinl(a);
            }
        p->twr = twr; p->twi = twi; p->wr = wr; p->wi = wi;
        __atomic_store_n(&p->N, N, __ATOMIC_RELEASE);
    }
    pthread_mutex_unlock(&g_lock);
    return p;
}

/* in-place DIF radix-2, natural in -> bit-reversed out, kernel exp(-2 pi i jk/M) */
static void fft_dif(double *restrict xr, double *restrict xi, int M, const plan_t *p) {
    for (int len = M; len >= 2; len >>= 1) {
        int half = len >> 1;
        const double *restrict wr = p->wr + half, *restrict wi = p->wi + half;
        for (int b = 0; b < M; b += len) {
            double *restrict ur = xr + b, *restrict ui = xi + b, *restrict vr = xr + b + half, *restrict vi = xi + b + half;
            for (int q = 0; q < half; q++) {
                double ar = ur[q], ai = ui[q], br = vr[q], bi = vi[q];
                double dr = ar - br, di = ai - bi;
                ur[q] = ar + br; ui[q] = ai + bi;
                vr[q] = dr * wr[q] - di * wi[q]; vi[q] = dr * wi[q] + di * wr[q];
            }
        }
    }
}
/* in-place DIT radix-2 inverse (unnormalised), bit-reversed in -> natural out, kernel exp(+2 pi i jk/M) */
static void fft_dit_inv(double *restrict xr, double *restrict xi, int M, const plan_t *p) {
    for (int len = 2; len <= M; len <<= 1) {
        int half = len >> 1;
        const double *restrict wr = p->wr + half, *restrict wi = p->wi + half;
        for (int b = 0; b < M; b += len) {
            double *restrict ur = xr + b, *restrict ui = xi + b, *restrict vr = xr + b + half, *restrict vi = xi + b + half;
            for (int q = 0; q < half; q++) {
                double cr = wr[q], ci = -wi[q];
                double br = vr[q] * cr - vi[q] * ci, bi = vr[q] * ci + vi[q] * cr;
                double ar = ur[q], ai = ui[q];
                ur[q] = ar + br; ui[q] = ai + bi;
                vr[q] = ar - br; vi[q] = ai - bi;
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
    double *restrict xr = out, *restrict xi = out + M;
    for (int j = 0; j < M; j++) {
        double re = to_signed_f64(poly[j], bits), im = to_signed_f64(poly[j + M], bits);
        xr[j] = re * p->twr[j] - im * p->twi[j];
        xi[j] = re * p->twi[j] + im * p->twr[j];
    }
    fft_dif(xr, xi, M, p);
}

/* fft/mod.rs:197-218 (convert_forward_torus: normalisation 2^-BITS) */
void orc_fft_forward_torus(int N, const uint64_t *poly, int bits, double *out) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    double norm = ldexp(1.0, -bits);
    double *restrict xr = out, *restrict xi = out + M;
    for (int j = 0; j < M; j++) {
        double re = to_signed_f64(poly[j], bits) * norm, im = to_signed_f64(poly[j + M], bits) * norm;
        xr[j] = re * p->twr[j] - im * p->twi[j];
        xi[j] = re * p->twi[j] + im * p->twr[j];
    }
    fft_dif(xr, xi, M, p);
}

/* commons/math/torus/mod.rs:72-78 from_torus, followed by the signed cast of fft/mod.rs:285-304 */
static inline uint64_t from_torus(double t, int bits) {
    /* rint (nearest-even) like the reference's AVX2/AVX-512 paths (fft/x86.rs:637-645); the scalar path's
     * round-half-away (torus/mod.rs:72-78) differs only on exact ties, which have measure zero */
    double fr = t - rint(t);
    fr *= ldexp(1.0, bits);
    fr = rint(fr);
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
    double *restrict xr = f, *restrict xi = f + M;
    fft_dit_inv(xr, xi, M, p);
    double inv = 1.0 / (double)M;
    for (int j = 0; j < M; j++) {
        double wr = p->twr[j], wi = -p->twi[j];
        double re = (xr[j] * wr - xi[j] * wi) * inv;
        double im = (xr[j] * wi + xi[j] * wr) * inv;
        poly[j] = (poly[j] + from_torus(re, bits)) & m;
        poly[j + M] = (poly[j + M] + from_torus(im, bits)) & m;
    }
}

/* pointwise complex multiply-accumulate o += g * d on structure-of-arrays Fourier polynomials (update_with_fmadd) */
void orc_fourier_mac(int M, double *restrict o, const double *restrict g, const double *restrict d) {
    double *restrict or_ = o, *restrict oi = o + M;
    const double *restrict gr = g, *restrict gi = g + M, *restrict dr = d, *restrict di = d + M;
    for (int f = 0; f < M; f++) {
        or_[f] += gr[f] * dr[f] - gi[f] * di[f];
        oi[f] += gr[f] * di[f] + gi[f] * dr[f];
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
            double r = fa[j] * fs[j] - fa[M + j] * fs[M + j];
            double i = fa[j] * fs[M + j] + fa[M + j] * fs[j];
            fa[j] = r; fa[M + j] = i;
        }
        fft_dit_inv(fa, fa + M, M, p);
        double inv = 1.0 / (double)M;
        for (int j = 0; j < M; j++) {
            double wr = p->twr[j], wi = -p->twi[j];
            double re = (fa[j] * wr - fa[M + j] * wi) * inv;
            double im = (fa[j] * wi + fa[M + j] * wr) * inv;
            out[j] += (uint64_t)(int64_t)llround(re) << sh;
            out[j + M] += (uint64_t)(int64_t)llround(im) << sh;
        }
    }
    for (int j = 0; j < N; j++) out[j] &= m;
    free(fs); free(fa); free(limb);
}

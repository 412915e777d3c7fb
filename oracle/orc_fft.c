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
 * that gcc vectorises the butterfly and pointwise loops (the reference gets its speed from pulp/AVX-512 intrinsics).
 *
 * The complex DFT is a radix-4 decimation-in-frequency forward / decimation-in-time inverse pair (one radix-2 stage first when
 * log2 M is odd), as concrete-fft's plans are radix-4/8/16 butterflies: half the passes over the data of a radix-2 transform.
 * The last stage (blocks of 4 consecutive points, twiddle 1) works on 16 points at a time through a 4x4 in-register transpose
 * and leaves its outputs transposed -- the Fourier-domain order is private to this file ("unordered" plan), forward and inverse
 * only have to agree, and pointwise products do not care. */
typedef double v4d __attribute__((vector_size(32), aligned(8)));
typedef long long v4i __attribute__((vector_size(32)));
typedef struct { int N, odd; double *twr, *twi;   /* twist e^{i pi j/N}, j < M */
                 double *w2r, *w2i;         /* radix-2 stage at len = M (odd log2 M only): exp(-2 pi i q/M), q < M/2 */
                 double *wq[32][6];         /* radix-4 stage of block length 2^s: w^q, w^2q, w^3q (re, im each), q < len/4, w = exp(-2 pi i/len) */
               } plan_t;
static plan_t g_plans[32];
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static double *alloc_d(size_t n) { return (double *)aligned_alloc(64, (sizeof(double) * (n + 8) + 63) & ~(size_t)63); }

static const plan_t *get_plan(int N) {
    int lg = 0; while ((1 << lg) < N) lg++;
    plan_t *p = &g_plans[lg];
    if (__atomic_load_n(&p->N, __ATOMIC_ACQUIRE) == N) return p;
    pthread_mutex_lock(&g_lock);
    if (p->N != N) {
        int M = N / 2, lgM = lg - 1;
        const long double PI = 3.14159265358979323846264338327950288L;
        double *twr = alloc_d(M), *twi = alloc_d(M);
        for (int j = 0; j < M; j++) {
            long double a = PI * (long double)j / (long double)N;
            twr[j] = (double)cosl(a); twi[j] = (double)sinl(a);
        }
        p->twr = twr; p->twi = twi;
        p->odd = lgM & 1;
        if (p->odd) {
            p->w2r = alloc_d(M / 2); p->w2i = alloc_d(M / 2);
            for (int q = 0; q < M / 2; q++) {
                long double a = -2.0L * PI * (long double)q / (long double)M;
                p->w2r[q] = (double)cosl(a); p->w2i[q] = (double)sinl(a);
            }
        }
        for (int s = 2; s <= lgM - p->odd; s += 2) {
            int len = 1 << s, q4 = len / 4;
            for (int e = 0; e < 3; e++) {
                double *wr = alloc_d(q4), *wi = alloc_d(q4);
                for (int q = 0; q < q4; q++) {
                    long double a = -2.0L * PI * (long double)((e + 1) * q) / (long double)len;
                    wr[q] = (double)cosl(a); wi[q] = (double)sinl(a);
                }
                p->wq[s][2 * e] = wr; p->wq[s][2 * e + 1] = wi;
            }
        }
        __atomic_store_n(&p->N, N, __ATOMIC_RELEASE);
    }
    pthread_mutex_unlock(&g_lock);
    return p;
}

/* one radix-4 DIF stage over every block of length len (len >= 16, so a quarter holds whole 4-wide vectors): sub-block e
 * receives the outputs k = e mod 4.  Written with explicit vectors: gcc's auto-vectoriser gives up on the 14 pointer streams. */
#define LD(p) (*(const v4d *)(p))
#define ST(p, v) (*(v4d *)(p) = (v))
static void r4_dif_stage(double *xr, double *xi, int M, int len, double *const *w) {
    const int q4 = len >> 2;
    for (int b = 0; b < M; b += len) {
        double *r0 = xr + b, *r1 = r0 + q4, *r2 = r1 + q4, *r3 = r2 + q4;
        double *i0 = xi + b, *i1 = i0 + q4, *i2 = i1 + q4, *i3 = i2 + q4;
        for (int q = 0; q < q4; q += 4) {
            const v4d a0r = LD(r0 + q), a0i = LD(i0 + q), a1r = LD(r1 + q), a1i = LD(i1 + q);
            const v4d a2r = LD(r2 + q), a2i = LD(i2 + q), a3r = LD(r3 + q), a3i = LD(i3 + q);
            const v4d t0r = a0r + a2r, t0i = a0i + a2i, t1r = a0r - a2r, t1i = a0i - a2i;
            const v4d t2r = a1r + a3r, t2i = a1i + a3i;
            const v4d t3r = a1i - a3i, t3i = a3r - a1r;              /* (a1 - a3) * (-i) */
            const v4d y1r = t1r + t3r, y1i = t1i + t3i, y2r = t0r - t2r, y2i = t0i - t2i, y3r = t1r - t3r, y3i = t1i - t3i;
            const v4d w1r = LD(w[0] + q), w1i = LD(w[1] + q), w2r = LD(w[2] + q), w2i = LD(w[3] + q), w3r = LD(w[4] + q), w3i = LD(w[5] + q);
            ST(r0 + q, t0r + t2r); ST(i0 + q, t0i + t2i);
            ST(r1 + q, y1r * w1r - y1i * w1i); ST(i1 + q, y1r * w1i + y1i * w1r);
            ST(r2 + q, y2r * w2r - y2i * w2i); ST(i2 + q, y2r * w2i + y2i * w2r);
            ST(r3 + q, y3r * w3r - y3i * w3i); ST(i3 + q, y3r * w3i + y3i * w3r);
        }
    }
}
/* its inverse (unnormalised: x4), conjugate twiddles */
static void r4_dit_stage(double *xr, double *xi, int M, int len, double *const *w) {
    const int q4 = len >> 2;
    for (int b = 0; b < M; b += len) {
        double *r0 = xr + b, *r1 = r0 + q4, *r2 = r1 + q4, *r3 = r2 + q4;
        double *i0 = xi + b, *i1 = i0 + q4, *i2 = i1 + q4, *i3 = i2 + q4;
        for (int q = 0; q < q4; q += 4) {
            const v4d w1r = LD(w[0] + q), w1i = LD(w[1] + q), w2r = LD(w[2] + q), w2i = LD(w[3] + q), w3r = LD(w[4] + q), w3i = LD(w[5] + q);
            const v4d y0r = LD(r0 + q), y0i = LD(i0 + q), y1r = LD(r1 + q), y1i = LD(i1 + q);
            const v4d y2r = LD(r2 + q), y2i = LD(i2 + q), y3r = LD(r3 + q), y3i = LD(i3 + q);
            const v4d z1r = y1r * w1r + y1i * w1i, z1i = y1i * w1r - y1r * w1i;
            const v4d z2r = y2r * w2r + y2i * w2i, z2i = y2i * w2r - y2r * w2i;
            const v4d z3r = y3r * w3r + y3i * w3i, z3i = y3i * w3r - y3r * w3i;
            const v4d t0r = y0r + z2r, t0i = y0i + z2i, t2r = y0r - z2r, t2i = y0i - z2i;
            const v4d t1r = z1r + z3r, t1i = z1i + z3i;
            const v4d u3r = z3i - z1i, u3i = z1r - z3r;              /* i * (z1 - z3) */
            ST(r0 + q, t0r + t1r); ST(i0 + q, t0i + t1i);
            ST(r2 + q, t0r - t1r); ST(i2 + q, t0i - t1i);
            ST(r1 + q, t2r + u3r); ST(i1 + q, t2i + u3i);
            ST(r3 + q, t2r - u3r); ST(i3 + q, t2i - u3i);
        }
    }
}

#define TRANSPOSE4(a, b, c, d)                                                      \
    do {                                                                            \
        v4d t0_ = __builtin_shuffle(a, b, (v4i){0, 4, 2, 6}), t1_ = __builtin_shuffle(a, b, (v4i){1, 5, 3, 7}); \
        v4d t2_ = __builtin_shuffle(c, d, (v4i){0, 4, 2, 6}), t3_ = __builtin_shuffle(c, d, (v4i){1, 5, 3, 7}); \
        a = __builtin_shuffle(t0_, t2_, (v4i){0, 1, 4, 5});                         \
        b = __builtin_shuffle(t1_, t3_, (v4i){0, 1, 4, 5});                         \
        c = __builtin_shuffle(t0_, t2_, (v4i){2, 3, 6, 7});                         \
        d = __builtin_shuffle(t1_, t3_, (v4i){2, 3, 6, 7});                         \
    } while (0)

/* last forward stage: 4-point DFTs of 4 consecutive points, 16 points at a time; output k of group g lands at 16 G + 4 k + g */
static void r4_dif_last(double *xr, double *xi, int M) {
    for (int b = 0; b < M; b += 16) {
        v4d ar = *(v4d *)(xr + b), br = *(v4d *)(xr + b + 4), cr = *(v4d *)(xr + b + 8), dr = *(v4d *)(xr + b + 12);
        v4d ai = *(v4d *)(xi + b), bi = *(v4d *)(xi + b + 4), ci = *(v4d *)(xi + b + 8), di = *(v4d *)(xi + b + 12);
        TRANSPOSE4(ar, br, cr, dr);
        TRANSPOSE4(ai, bi, ci, di);
        v4d t0r = ar + cr, t0i = ai + ci, t1r = ar - cr, t1i = ai - ci, t2r = br + dr, t2i = bi + di, t3r = bi - di, t3i = dr - br;
        *(v4d *)(xr + b) = t0r + t2r;      *(v4d *)(xi + b) = t0i + t2i;
        *(v4d *)(xr + b + 4) = t1r + t3r;  *(v4d *)(xi + b + 4) = t1i + t3i;
        *(v4d *)(xr + b + 8) = t0r - t2r;  *(v4d *)(xi + b + 8) = t0i - t2i;
        *(v4d *)(xr + b + 12) = t1r - t3r; *(v4d *)(xi + b + 12) = t1i - t3i;
    }
}
static void r4_dit_first(double *xr, double *xi, int M) {
    for (int b = 0; b < M; b += 16) {
        v4d y0r = *(v4d *)(xr + b), y1r = *(v4d *)(xr + b + 4), y2r = *(v4d *)(xr + b + 8), y3r = *(v4d *)(xr + b + 12);
        v4d y0i = *(v4d *)(xi + b), y1i = *(v4d *)(xi + b + 4), y2i = *(v4d *)(xi + b + 8), y3i = *(v4d *)(xi + b + 12);
        v4d t0r = y0r + y2r, t0i = y0i + y2i, t2r = y0r - y2r, t2i = y0i - y2i, t1r = y1r + y3r, t1i = y1i + y3i;
        v4d u3r = y3i - y1i, u3i = y1r - y3r;
        v4d ar = t0r + t1r, ai = t0i + t1i, cr = t0r - t1r, ci = t0i - t1i, br = t2r + u3r, bi = t2i + u3i, dr = t2r - u3r, di = t2i - u3i;
        TRANSPOSE4(ar, br, cr, dr);
        TRANSPOSE4(ai, bi, ci, di);
        *(v4d *)(xr + b) = ar; *(v4d *)(xr + b + 4) = br; *(v4d *)(xr + b + 8) = cr; *(v4d *)(xr + b + 12) = dr;
        *(v4d *)(xi + b) = ai; *(v4d *)(xi + b + 4) = bi; *(v4d *)(xi + b + 8) = ci; *(v4d *)(xi + b + 12) = di;
    }
}

/* in-place forward DFT, natural in -> this file's private order out, kernel exp(-2 pi i jk/M); M >= 16 */
static void fft_dif(double *restrict xr, double *restrict xi, int M, const plan_t *p) {
    int len = M;
    if (p->odd) {
        const int half = M >> 1;
        const double *restrict wr = p->w2r, *restrict wi = p->w2i;
        double *restrict ur = xr, *restrict ui = xi, *restrict vr = xr + half, *restrict vi = xi + half;
        for (int q = 0; q < half; q++) {
            double ar = ur[q], ai = ui[q], br = vr[q], bi = vi[q];
            double dr = ar - br, di = ai - bi;
            ur[q] = ar + br; ui[q] = ai + bi;
            vr[q] = dr * wr[q] - di * wi[q]; vi[q] = dr * wi[q] + di * wr[q];
        }
        len = half;
    }
    int s = 0; while ((1 << s) < len) s++;
    for (; len > 4; len >>= 2, s -= 2) r4_dif_stage(xr, xi, M, len, p->wq[s]);
    r4_dif_last(xr, xi, M);
}
/* in-place inverse (unnormalised: x M), this file's private order in -> natural out, kernel exp(+2 pi i jk/M) */
static void fft_dit_inv(double *restrict xr, double *restrict xi, int M, const plan_t *p) {
    const int top = p->odd ? M >> 1 : M;
    r4_dit_first(xr, xi, M);
    int s = 4;
    for (int len = 16; len <= top; len <<= 2, s += 2) r4_dit_stage(xr, xi, M, len, p->wq[s]);
    if (p->odd) {
        const int half = M >> 1;
        const double *restrict wr = p->w2r, *restrict wi = p->w2i;
        double *restrict ur = xr, *restrict ui = xi, *restrict vr = xr + half, *restrict vi = xi + half;
        for (int q = 0; q < half; q++) {
            double br = vr[q] * wr[q] + vi[q] * wi[q], bi = vi[q] * wr[q] - vr[q] * wi[q];
            double ar = ur[q], ai = ui[q];
            ur[q] = ar + br; ui[q] = ai + bi;
            vr[q] = ar - br; vi[q] = ai - bi;
        }
    }
}

/* into_signed().cast_into() (fft/mod.rs:228-229) for a whole polynomial, written so that gcc vectorises it without AVX-512:
 * v = hi * 2^32 + lo with hi signed and lo unsigned 32-bit halves, both conversions exact, ONE rounding in the final addition
 * -- the same correctly rounded value as the scalar i64 -> f64 conversion. */
static void signed_to_f64(const uint64_t *restrict poly, int n, int bits, double scale, double *restrict out) {
    if (bits == 64) {
        for (int j = 0; j < n; j++) {
            const int32_t hi = (int32_t)(poly[j] >> 32);
            const int32_t lo = (int32_t)((uint32_t)poly[j] ^ 0x80000000u);          /* lo - 2^31 as a signed value */
            out[j] = ((double)hi * 4294967296.0 + ((double)lo + 2147483648.0)) * scale;
        }
    } else {
        for (int j = 0; j < n; j++) out[j] = (double)(int32_t)(uint32_t)poly[j] * scale;
    }
}

/* fft/mod.rs:220-239 (convert_forward_integer) + :496-515 */
static void forward_scaled(int N, const uint64_t *poly, int bits, double scale, double *out) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    double *restrict xr = out, *restrict xi = out + M;
    const double *restrict twr = p->twr, *restrict twi = p->twi;
    signed_to_f64(poly, N, bits, scale, out);
    for (int j = 0; j < M; j++) {
        double re = xr[j], im = xi[j];
        xr[j] = re * twr[j] - im * twi[j];
        xi[j] = re * twi[j] + im * twr[j];
    }
    fft_dif(xr, xi, M, p);
}
void orc_fft_forward_integer(int N, const uint64_t *poly, int bits, double *out) { forward_scaled(N, poly, bits, 1.0, out); }

/* fft/mod.rs:197-218 (convert_forward_torus: normalisation 2^-BITS, an exact scaling) */
void orc_fft_forward_torus(int N, const uint64_t *poly, int bits, double *out) { forward_scaled(N, poly, bits, ldexp(1.0, -bits), out); }

/* commons/math/torus/mod.rs:72-78 from_torus, followed by the signed cast of fft/mod.rs:285-304, for a whole array, added
 * (wrapping) into poly.  rint (nearest-even) like the reference's AVX2/AVX-512 paths (fft/x86.rs:637-645); the scalar
 * path's round-half-away (torus/mod.rs:72-78) differs only on exact ties, which have measure zero.  The f64 -> i64 cast is
 * done on exact 32-bit halves so that it vectorises; Rust's `as i64` saturates, which is only reachable at fr == 2^63. */
static void add_from_torus(uint64_t *restrict poly, const double *restrict t, int n, int bits) {
    if (bits == 64) {
        for (int j = 0; j < n; j++) {
            double fr = t[j] - rint(t[j]);
            fr = rint(fr * 18446744073709551616.0);                    /* integer-valued, |fr| <= 2^63 */
            const double hi = floor(fr * (1.0 / 4294967296.0));        /* exact: a power-of-two scaling of an integer */
            const double lo = fr - hi * 4294967296.0;                  /* exact, in [0, 2^32) */
            const uint32_t h = (uint32_t)(int32_t)hi;
            const uint32_t l = (uint32_t)(int32_t)(lo - 2147483648.0) ^ 0x80000000u;
            uint64_t v = ((uint64_t)h << 32) | l;
            v = fr >= 9223372036854775808.0 ? (uint64_t)INT64_MAX : v;
            poly[j] += v;
        }
    } else {
        for (int j = 0; j < n; j++) {
            double fr = t[j] - rint(t[j]);
            fr = rint(fr * 4294967296.0);
            const int64_t v = (int64_t)fr;                             /* |fr| <= 2^31: no saturation */
            poly[j] = (poly[j] + (uint64_t)v) & 0xFFFFFFFFull;
        }
    }
}

/* fft/mod.rs:285-326,539-557 : inverse FFT, * conj(twist) / M, round to torus, wrapping add */
void orc_fft_add_backward_torus(int N, uint64_t *poly, int bits, double *f) {
    const plan_t *p = get_plan(N);
    int M = N / 2;
    double *restrict xr = f, *restrict xi = f + M;
    const double *restrict twr = p->twr, *restrict twi = p->twi;
    fft_dit_inv(xr, xi, M, p);
    double inv = 1.0 / (double)M;
    for (int j = 0; j < M; j++) {
        double wr = twr[j], wi = -twi[j];
        double re = (xr[j] * wr - xi[j] * wi) * inv;
        double im = (xr[j] * wi + xi[j] * wr) * inv;
        xr[j] = re; xi[j] = im;
    }
    add_from_torus(poly, f, N, bits);
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

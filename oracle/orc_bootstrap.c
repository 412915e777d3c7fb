/* orc_bootstrap.c -- external product, blind rotation, PBS, batched KS+PBS
 * (TEST INFRASTRUCTURE, see tfhe_oracle.h). */
#include "tfhe_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

static inline uint64_t maskw(int bits) { return bits == 64 ? ~0ULL : ((1ULL << bits) - 1); }

/* generic "run fn(i) for i in [0,count) over nthreads" helper (rayon par_iter stand-in) */
typedef struct { void (*fn)(void *, int); void *ctx; int count; int next; pthread_mutex_t mu; } pool_t;
static void *pool_worker(void *arg) {
    pool_t *p = (pool_t *)arg;
    for (;;) {
        pthread_mutex_lock(&p->mu);
        int i = p->next++;
        pthread_mutex_unlock(&p->mu);
        if (i >= p->count) break;
        p->fn(p->ctx, i);
    }
    return NULL;
}
void orc_parallel_for(void (*fn)(void *, int), void *ctx, int count, int nthreads) {
    if (nthreads <= 1 || count <= 1) { for (int i = 0; i < count; i++) fn(ctx, i); return; }
    pool_t p; p.fn = fn; p.ctx = ctx; p.count = count; p.next = 0;
    pthread_mutex_init(&p.mu, NULL);
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, pool_worker, &p);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    pthread_mutex_destroy(&p.mu);
}

/* ---- BSK conversion: fft64/crypto/bootstrap.rs:197-209 -> fft/mod.rs:719-764 ---- */
typedef struct { const orc_params *p; const uint64_t *bsk; double *f; } conv_ctx;
static void conv_one(void *c, int poly) {
    conv_ctx *x = (conv_ctx *)c;
    int N = x->p->N;
    orc_fft_forward_torus(N, x->bsk + (size_t)poly * N, x->p->bits, x->f + (size_t)poly * N);
}
void orc_bsk_to_fourier(const orc_params *p, const uint64_t *bsk, double *fbsk, int nthreads) {
    conv_ctx c = { p, bsk, fbsk };
    int polys = p->n * p->pbs_level * (p->k + 1) * (p->k + 1);
    orc_parallel_for(conv_one, &c, polys, nthreads);
}

void orc_fourier_mac(int M, double *o, const double *g, const double *d);

/* per-thread scratch (the reference's PodStack / ComputationBuffers, computation_buffers.rs:8-29) */
typedef struct { size_t cap; uint64_t *states, *digits, *tmp; double *fourier, *outf; } scratch_t;
static __thread scratch_t g_scr;
static scratch_t *get_scratch(size_t tot, int N) {
    if (g_scr.cap < tot) {
        free(g_scr.states); free(g_scr.digits); free(g_scr.tmp); free(g_scr.fourier); free(g_scr.outf);
        g_scr.states = (uint64_t *)aligned_alloc(64, sizeof(uint64_t) * tot);
        g_scr.tmp = (uint64_t *)aligned_alloc(64, sizeof(uint64_t) * tot);
        g_scr.digits = (uint64_t *)aligned_alloc(64, sizeof(uint64_t) * N);
        g_scr.fourier = (double *)aligned_alloc(64, sizeof(double) * N);
        g_scr.outf = (double *)aligned_alloc(64, sizeof(double) * tot);
        g_scr.cap = tot;
    }
    return &g_scr;
}

/* fft64/crypto/ggsw.rs:477-598.  ggsw_f: [level 1..l][row][col][N doubles: re | im] (levels stored 1->l, consumed l->1).
 * The integer loops are written branch-free over whole polynomials so that gcc vectorises them (the reference's are SIMD too:
 * fft64/math/decomposition.rs:27-86 works on slices). */
void orc_add_external_product(const orc_params *p, uint64_t *acc, const double *ggsw_f, const uint64_t *glwe) {
    const int N = p->N, M = N / 2, k1 = p->k + 1, L = p->pbs_level, b = p->pbs_base_log, bits = p->bits;
    const uint64_t m = maskw(bits), mb = (1ULL << b) - 1;
    const size_t tot = (size_t)k1 * N;
    scratch_t *sc = get_scratch(tot, N);
    uint64_t *restrict states = sc->states, *restrict digits = sc->digits;
    double *fourier = sc->fourier, *outf = sc->outf;
    memset(outf, 0, sizeof(double) * tot);
    /* SignedDecomposer::closest_representable (decomposer.rs:99-119) then the iterator's initial shift (iter.rs:37-50) */
    const int shift = bits - b * L, rshift = shift - 1;
    for (size_t t = 0; t < tot; t++) {
        uint64_t res = (glwe[t] & m) >> rshift;
        res += 1;
        res &= ~1ULL;
        states[t] = ((res << rshift) & m) >> shift;
    }
    for (int lvl = L; lvl >= 1; lvl--) {
        const double *mat = ggsw_f + (size_t)(lvl - 1) * k1 * k1 * N;
        for (int r = 0; r < k1; r++) {
            /* collect_next_term: ggsw.rs:601-613 ; decompose_one_level: decomposition.rs:79-86 */
            uint64_t *restrict st = states + (size_t)r * N;
            for (int j = 0; j < N; j++) {
                const uint64_t res = st[j] & mb;
                const uint64_t s1 = st[j] >> b;
                uint64_t carry = (((res - 1) & m) | s1) & res;
                carry >>= (b - 1);
                st[j] = (s1 + carry) & m;
                digits[j] = (res - (carry << b)) & m;
            }
            orc_fft_forward_integer(N, digits, bits, fourier);
            /* update_with_fmadd: ggsw.rs:616-697 */
            const double *row = mat + (size_t)r * k1 * N;
            for (int j = 0; j < k1; j++) orc_fourier_mac(M, outf + (size_t)j * N, row + (size_t)j * N, fourier);
        }
    }
    for (int j = 0; j < k1; j++) orc_fft_add_backward_torus(N, acc + (size_t)j * N, bits, outf + (size_t)j * N);
}

/* fft64/crypto/bootstrap.rs:243-344 */
void orc_blind_rotate(const orc_params *p, const double *fbsk, const uint64_t *lwe, uint64_t *acc) {
    int N = p->N, k1 = p->k + 1, n = p->n, bits = p->bits;
    int log2N = 0; while ((1 << log2N) < N) log2N++;
    size_t tot = (size_t)k1 * N;
    uint64_t *tmp = get_scratch(tot, N)->tmp;
    uint64_t bhat = orc_modswitch(lwe[n], log2N, bits);
    memcpy(tmp, acc, sizeof(uint64_t) * tot);
    for (int j = 0; j < k1; j++) orc_monomial_div(acc + (size_t)j * N, tmp + (size_t)j * N, N, bhat, bits);
    size_t ggsw_sz = (size_t)p->pbs_level * k1 * k1 * N; /* doubles */
    for (int i = 0; i < n; i++) {
        if ((lwe[i] & maskw(bits)) == 0) continue; /* bootstrap.rs:285 */
        uint64_t ahat = orc_modswitch(lwe[i], log2N, bits);
        for (int j = 0; j < k1; j++) orc_monomial_mul_sub(tmp + (size_t)j * N, acc + (size_t)j * N, N, ahat, bits);
        orc_add_external_product(p, acc, fbsk + (size_t)i * ggsw_sz, tmp);
    }
}

/* bootstrap.rs:346-380 */
void orc_pbs(const orc_params *p, const double *fbsk, const uint64_t *lwe_in, const uint64_t *lut, uint64_t *lwe_out) {
    size_t tot = (size_t)(p->k + 1) * p->N;
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * tot);
    memcpy(acc, lut, sizeof(uint64_t) * tot);
    orc_blind_rotate(p, fbsk, lwe_in, acc);
    orc_sample_extract0(lwe_out, acc, p->k, p->N, p->bits);
    free(acc);
}

/* ---- multi-bit PBS: algorithms/lwe_multi_bit_programmable_bootstrapping.rs -------------------------------------------- */
/* prepare_multi_bit_ggsw_mem_optimized :18-85.  group_f: the 2^G Fourier GGSWs of one group; out = GGSW_0 +
 * sum_{idx>=1} GGSW_idx (.) FFT(X^{modswitch(sum of the selected mask elements)}).  The reference obtains the monomial's
 * transform from Plan::fwd_monomial times a twist factor (fft/mod.rs:408-445); that is the forward transform of the
 * monomial, which is what is computed here. */
static void multibit_prepare_ggsw(const orc_params *p, int G, const double *group_f, const uint64_t *mask_elems, double *out,
                                  uint64_t *mono, double *mono_f) {
    int N = p->N, M = N / 2, k1 = p->k + 1, bits = p->bits;
    int log2N = 0; while ((1 << log2N) < N) log2N++;
    size_t npoly = (size_t)p->pbs_level * k1 * k1, ggsw_sz = npoly * N;
    memcpy(out, group_f, sizeof(double) * ggsw_sz);
    for (int idx = 1; idx < (1 << G); idx++) {
        uint64_t deg = 0;
        for (int b = 0; b < G; b++) if ((idx >> (G - (b + 1))) & 1) deg += mask_elems[b];
        uint64_t sw = orc_modswitch(deg & maskw(bits), log2N, bits);       /* in [0, 2N] */
        memset(mono, 0, sizeof(uint64_t) * N);
        mono[sw % N] = ((sw / N) & 1) ? maskw(bits) : 1;                    /* X^N = -1 */
        orc_fft_forward_integer(N, mono, bits, mono_f);
        const double *g = group_f + (size_t)idx * ggsw_sz;
        for (size_t q = 0; q < npoly; q++) orc_fourier_mac(M, out + q * N, g + q * N, mono_f);
    }
}

/* multi_bit_deterministic_blind_rotate_assign :548-… (groups consumed in ascending order; the non-deterministic variant
 * :295-546 consumes them in whatever order its producer threads deliver, which changes only f64 rounding).
 * mb_fbsk: [n/G][2^G] Fourier GGSWs. */
void orc_multibit_blind_rotate(const orc_params *p, int G, const double *mb_fbsk, const uint64_t *lwe, uint64_t *acc) {
    int N = p->N, k1 = p->k + 1, n = p->n, bits = p->bits;
    int log2N = 0; while ((1 << log2N) < N) log2N++;
    size_t tot = (size_t)k1 * N;
    size_t ggsw_sz = (size_t)p->pbs_level * k1 * k1 * N;
    uint64_t *src = (uint64_t *)malloc(sizeof(uint64_t) * tot), *mono = (uint64_t *)malloc(sizeof(uint64_t) * N);
    double *buf = (double *)malloc(sizeof(double) * ggsw_sz), *mono_f = (double *)malloc(sizeof(double) * N);
    uint64_t bhat = orc_modswitch(lwe[n], log2N, bits);
    for (int j = 0; j < k1; j++) orc_monomial_div(src + (size_t)j * N, acc + (size_t)j * N, N, bhat, bits);
    for (int gi = 0; gi < n / G; gi++) {
        multibit_prepare_ggsw(p, G, mb_fbsk + (size_t)gi * (1 << G) * ggsw_sz, lwe + (size_t)gi * G, buf, mono, mono_f);
        memset(acc, 0, sizeof(uint64_t) * tot);                   /* dst_ct.as_mut().fill(ZERO) */
        orc_add_external_product(p, acc, buf, src);
        memcpy(src, acc, sizeof(uint64_t) * tot);
    }
    memcpy(acc, src, sizeof(uint64_t) * tot);
    free(src); free(mono); free(buf); free(mono_f);
}

/* multi_bit_(deterministic_)programmable_bootstrap_lwe_ciphertext :1035-1224 */
void orc_multibit_pbs(const orc_params *p, int G, const double *mb_fbsk, const uint64_t *lwe_in, const uint64_t *lut, uint64_t *lwe_out) {
    size_t tot = (size_t)(p->k + 1) * p->N;
    uint64_t *acc = (uint64_t *)malloc(sizeof(uint64_t) * tot);
    memcpy(acc, lut, sizeof(uint64_t) * tot);
    orc_multibit_blind_rotate(p, G, mb_fbsk, lwe_in, acc);
    orc_sample_extract0(lwe_out, acc, p->k, p->N, p->bits);
    free(acc);
}

typedef struct {
    const orc_params *p; const uint64_t *ksk; const double *fbsk; const uint64_t *in; const uint64_t *luts;
    const uint32_t *lut_idx; uint64_t *out; int mode; /* 0 ks->pbs, 1 pbs->ks, 2 pbs, 3 blind rotate only */
    int grouping;     /* 0: classic bootstrap key; G >= 2: multi-bit key with grouping factor G */
} batch_ctx;
static void any_pbs(const batch_ctx *x, const uint64_t *in, const uint64_t *lut, uint64_t *out) {
    if (x->grouping) orc_multibit_pbs(x->p, x->grouping, x->fbsk, in, lut, out);
    else orc_pbs(x->p, x->fbsk, in, lut, out);
}
static void batch_one(void *c, int i) {
    batch_ctx *x = (batch_ctx *)c;
    const orc_params *p = x->p;
    size_t big = (size_t)p->k * p->N + 1, small = (size_t)p->n + 1, glwe = (size_t)(p->k + 1) * p->N;
    const uint64_t *lut = x->luts + (x->lut_idx ? (size_t)x->lut_idx[i] * glwe : 0);
    if (x->mode == 0) {
        uint64_t *mid = (uint64_t *)malloc(sizeof(uint64_t) * small);
        orc_keyswitch(p, x->ksk, x->in + (size_t)i * big, mid);
        any_pbs(x, mid, lut, x->out + (size_t)i * big);
        free(mid);
    } else if (x->mode == 1) {
        uint64_t *mid = (uint64_t *)malloc(sizeof(uint64_t) * big);
        any_pbs(x, x->in + (size_t)i * small, lut, mid);
        orc_keyswitch(p, x->ksk, mid, x->out + (size_t)i * small);
        free(mid);
    } else if (x->mode == 2) {
        any_pbs(x, x->in + (size_t)i * small, lut, x->out + (size_t)i * big);
    } else {
        uint64_t *acc = x->out + (size_t)i * glwe;
        memcpy(acc, lut, sizeof(uint64_t) * glwe);
        if (x->grouping) orc_multibit_blind_rotate(p, x->grouping, x->fbsk, x->in + (size_t)i * small, acc);
        else orc_blind_rotate(p, x->fbsk, x->in + (size_t)i * small, acc);
    }
}
void orc_ks_pbs_batch(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in,
                      const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads) {
    batch_ctx c = { p, ksk, fbsk, in, luts, lut_idx, out, 0, 0 };
    orc_parallel_for(batch_one, &c, batch, nthreads);
}
void orc_pbs_ks_batch(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in,
                      const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads) {
    batch_ctx c = { p, ksk, fbsk, in, luts, lut_idx, out, 1, 0 };
    orc_parallel_for(batch_one, &c, batch, nthreads);
}
void orc_pbs_batch(const orc_params *p, const double *fbsk, const uint64_t *in, const uint64_t *luts,
                   const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads, int sample_extract) {
    batch_ctx c = { p, NULL, fbsk, in, luts, lut_idx, out, sample_extract ? 2 : 3, 0 };
    orc_parallel_for(batch_one, &c, batch, nthreads);
}

/* shortint ServerKey with ShortintBootstrappingKey::MultiBit (shortint/server_key/mod.rs:829-851), batched.
 * mode: 0 KS->PBS, 2 PBS, 3 blind rotation without sample extract */
void orc_multibit_batch(const orc_params *p, int grouping, int mode, const uint64_t *ksk, const double *mb_fbsk, const uint64_t *in,
                        const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads) {
    batch_ctx c = { p, ksk, mb_fbsk, in, luts, lut_idx, out, mode, grouping };
    orc_parallel_for(batch_one, &c, batch, nthreads);
}

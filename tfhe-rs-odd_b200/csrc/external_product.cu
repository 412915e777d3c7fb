// external_product.cu -- stand-alone GGSW (x) GLWE external product and CMUX over a batch.
//
// Reference: add_external_product_assign (core_crypto/algorithms/lwe_programmable_bootstrapping.rs:309-475 ->
// fft_impl/fft64/crypto/ggsw.rs:477-598) and cmux_assign (:552-765 -> ggsw.rs:766-777):
//     out <- out + ggsw (x) glwe                       cmux:  ct1 <- ct1 - ct0 ;  ct0 <- ct0 + ggsw (x) ct1
// These are the body of one blind-rotation iteration without the rotate-and-subtract, exported so that ONE external product
// can be pinned at FFT precision on the device (tests/test_external_product.py).  The stages are the shared ones of br_core.cuh / br_big.cuh
// (signed decomposition, forward team FFT, Fourier multiply-accumulate in the reference's summation order, inverse FFT,
// torus rounding, wrapping add); one CTA per ciphertext, operands read from and written to global memory.
#include "br_big.cuh"
#include "kernels.h"

namespace b200tfhe {

template <typename T, int R, int K, int L, int BL> struct EPConfig {
    typedef BRCore<T, R, K, L, BL> C;
    static constexpr int K1 = K + 1;
    static constexpr int CTS = (K1 * R) % 32 == 0 ? 1 : 2;      // ciphertexts per CTA: teams must fill whole warps (boolean: 3 x 16 threads)
    static constexpr int THREADS = CTS * K1 * R;
    static_assert(THREADS % 32 == 0, "teams must fill whole warps");
    static constexpr size_t FBUF_BYTES = (size_t)K1 * L * C::REGION * sizeof(cplx);
    static constexpr size_t POLY_BYTES = (size_t)K1 * C::N * sizeof(T);
    static constexpr size_t CT_BYTES = FBUF_BYTES + 2 * POLY_BYTES;
    static constexpr size_t TW_BYTES = sizeof(TwTables<R>);
    static constexpr size_t SMEM = CTS * CT_BYTES + TW_BYTES;
};

// grid = ceil(batch / CTS).  ciphertext b:  cmux == 0:  out[b] += GGSW[idx[b]] (x) src[b]
//                        cmux != 0:  src[b] <- src[b] - out[b] ;  out[b] += GGSW[idx[b]] (x) src[b]      (out = ct0, src = ct1)
template <typename T, int R, int K, int L, int BL>
__global__ void __launch_bounds__(EPConfig<T, R, K, L, BL>::THREADS)
external_product_kernel(T *__restrict__ out, T *__restrict__ src, const cplx *__restrict__ ggsw_list, const uint32_t *__restrict__ idx,
                        int cmux, int batch, const cplx *__restrict__ tw_fwd_, const cplx *__restrict__ tw_inv_) {
    typedef EPConfig<T, R, K, L, BL> CFG;
    typedef typename CFG::C C;
    constexpr int K1 = K + 1, N = C::N, REGION = C::REGION;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = threadIdx.x / (K1 * R), r = (threadIdx.x / R) % K1, lane = threadIdx.x % R;
    size_t ct = (size_t)blockIdx.x * CFG::CTS + q;
    const bool live = ct < (size_t)batch;      // a ghost ciphertext (odd batch, CTS = 2) recomputes the last one and writes nothing
    if (!live) ct = (size_t)batch - 1;
    const size_t glwe = (size_t)K1 * N;

    unsigned char *ct_smem = smem_raw + (size_t)q * CFG::CT_BYTES;
    cplx *fbuf_ct = reinterpret_cast<cplx *>(ct_smem);
    cplx *fbuf_r = fbuf_ct + (size_t)r * L * REGION;
    T *acc_r = reinterpret_cast<T *>(ct_smem + CFG::FBUF_BYTES) + (size_t)r * N;
    T *in_r = reinterpret_cast<T *>(ct_smem + CFG::FBUF_BYTES + CFG::POLY_BYTES) + (size_t)r * N;
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + CFG::CTS * CFG::CT_BYTES);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int q = threadIdx.x; q < TW_HALF; q += CFG::THREADS) {
        tw_s[q] = tw_fwd_[q];
        tw_s[TW_HALF + q] = tw_inv_[q];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    T *o_r = out + ct * glwe + (size_t)r * N;
    T *s_r = src + ct * glwe + (size_t)r * N;
    for (int c = lane; c < N; c += R) {
        const T a = o_r[c];
        T d = s_r[c];
        if (cmux) {
            d -= a;                                  // ggsw.rs:773: ct1 -= ct0
            if (live) s_r[c] = d;
        }
        acc_r[c] = a;
        in_r[c] = d;
    }
    __syncthreads();

    const cplx *g = ggsw_list + (size_t)(idx ? idx[ct] : 0) * C::GGSW_CPLX;
    plain_digits_fwd_pass1<T, R, K, L, BL>(lane, in_r, fbuf_r, twf);
    __syncwarp();
    cplx F[R];
    C::template mac_prefetch_part<0>(lane, r, g, F);
#pragma unroll 1
    for (int t = 0; t < L; t++) {
        cplx x[R];
        C::fwd_pass2_load(lane, x, fbuf_r + t * REGION);
        __syncwarp();
        C::fwd_pass2_store(lane, x, fbuf_r + t * REGION);
    }
    if constexpr (R > C::CH) C::template mac_prefetch_part<1>(lane, r, g, F);
    __syncthreads();
    C::mac_finish(lane, r, g, fbuf_ct, F);
    __syncthreads();
    C::inv_pass1(lane, F, fbuf_r, twi);
    __syncwarp();
    C::inv_pass2_accumulate(lane, fbuf_r, acc_r);
    __syncwarp();
    if (live)
        for (int c = lane; c < N; c += R) o_r[c] = acc_r[c];
}

// N = 8192: 256 threads per ciphertext; the output polynomial is the outer loop (the forward transforms are redone per output
// polynomial: this entry point exists for parity checks, not for throughput), so only one Fourier accumulator is live.
template <typename T, int K, int L, int BL>
__global__ void __launch_bounds__(256, 1)
external_product_big_kernel(T *__restrict__ out, T *__restrict__ src, const cplx *__restrict__ ggsw_list, const uint32_t *__restrict__ idx,
                            int cmux, const BigTables *__restrict__ tb) {
    typedef BigCore<T, K, L, BL> C;
    constexpr int K1 = K + 1, N = C::N, P = C::P;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx *S1 = reinterpret_cast<cplx *>(smem_raw);
    const int tid = threadIdx.x;
    const size_t ct = blockIdx.x, glwe = (size_t)K1 * N;
    T *o = out + ct * glwe, *s = src + ct * glwe;
    if (cmux) {
        for (int c = tid; c < K1 * N; c += 256) s[c] -= o[c];
        __syncthreads();           // the CTA's own global writes are visible to it after the barrier
    }
    const cplx *ggsw = ggsw_list + (size_t)(idx ? idx[ct] : 0) * C::GGSW_CPLX;
#pragma unroll 1
    for (int j = 0; j < K1; j++) {
        cplx f[P];
        static_for<0, P>([&](auto p) { f[decltype(p)::value] = cplx{0.0, 0.0}; });
        // level L down to 1 (t ascending), rows ascending: the reference's summation order (ggsw.rs:524-577)
#pragma unroll 1
        for (int st = 0; st < L * K1; st++) {
            const int t = st / K1, r = st % K1, lvl = L - 1 - t;
            cplx x[P];
            C::digits_plain_rt(tid, s + (size_t)r * N, t, x);
            __syncthreads();
            C::fwdA(tid, x, S1, tb->ta, tb->tq);
            __syncthreads();
            C::fwdB_load(tid, x, S1);
            __syncwarp();
            C::fwdB_pass1(tid, x, S1, tb);
            __syncwarp();
            C::fwdB_pass2(tid, x, S1);
            const cplx *gbase = ggsw + ((size_t)(lvl * K1 + r) * K1 + j) * C::M + tid;
            cplx gv[P];
            static_for<0, P>([&](auto p) { gv[decltype(p)::value] = ldg_stream(gbase + decltype(p)::value * 256); });
            C::mac_one_poly(gv, x, f);
        }
        cplx x[P];
        static_for<0, P>([&](auto pp) { x[decltype(pp)::value] = f[brev(decltype(pp)::value, C::LOGP)]; });
        __syncthreads();
        C::invA(tid, x, S1, tb);
        __syncwarp();
        C::invB_load(tid, x, S1);
        __syncwarp();
        C::invB_store(tid, x, S1);
        __syncthreads();
        C::invC(tid, x, S1, tb->ta, tb->tq);
        C::accumulate(tid, x, o + (size_t)j * N);
    }
}

// every parameter set the blind rotation is built for (BR_INSTANCES of blind_rotate.cu) + N = 8192 (BIG_INSTANCES)
#define EP_INSTANCES(X)           \
    X(uint64_t, 32, 1, 2, 15)     \
    X(uint64_t, 32, 1, 3, 5)      \
    X(uint64_t, 16, 3, 2, 12)     \
    X(uint64_t, 16, 3, 4, 3)      \
    X(uint64_t, 16, 3, 3, 9)      \
    X(uint64_t, 16, 3, 5, 3)      \
    X(uint64_t, 32, 1, 1, 23)     \
    X(uint32_t, 16, 2, 3, 6)      \
    X(uint64_t, 16, 3, 1, 18)     \
    X(uint64_t, 16, 1, 2, 12)
#define EP_BIG_INSTANCES(X) X(uint64_t, 1, 2, 15)

int external_product_supported(const Params &p) {
#define X(T, R, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return R;
    EP_INSTANCES(X)
#undef X
#define X(T, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 8192 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return 256;
    EP_BIG_INSTANCES(X)
#undef X
    return 0;
}

int external_product_launch(const Params &p, void *out, void *src, const void *ggsw_list, const uint32_t *idx, int cmux, int batch,
                            const void *tw_fwd, const void *tw_inv, cudaStream_t st) {
    if (batch <= 0) return 0;
#define X(T, R, K, L, BL)                                                                                                      \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) {              \
        typedef EPConfig<T, R, K, L, BL> CFG;                                                                                  \
        auto kern = external_product_kernel<T, R, K, L, BL>;                                                                    \
        if (cudaError_t e = ensure_dynamic_smem((const void *)kern, CFG::SMEM)) return (int)e;                                  \
        kern<<<(unsigned)((batch + CFG::CTS - 1) / CFG::CTS), CFG::THREADS, CFG::SMEM, st>>>(                                    \
            (T *)out, (T *)src, (const cplx *)ggsw_list, idx, cmux, batch, (const cplx *)tw_fwd, (const cplx *)tw_inv);          \
        return (int)cudaGetLastError();                                                                                        \
    }
    EP_INSTANCES(X)
#undef X
#define X(T, K, L, BL)                                                                                                         \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 8192 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) {                    \
        auto kern = external_product_big_kernel<T, K, L, BL>;                                                                   \
        if (cudaError_t e = ensure_dynamic_smem((const void *)kern, 65536)) return (int)e;                                      \
        kern<<<(unsigned)batch, 256, 65536, st>>>((T *)out, (T *)src, (const cplx *)ggsw_list, idx, cmux, (const BigTables *)tw_fwd); \
        return (int)cudaGetLastError();                                                                                        \
    }
    EP_BIG_INSTANCES(X)
#undef X
    return -1;
}

}  // namespace b200tfhe

// blind_rotate_pair.cu -- blind rotation for the (k = 3, N = 512, l = 1) shape of PARAM_MESSAGE_1_CARRY_1 (Trivium's set):
// SIX ciphertexts per CTA, two per warp.
//
// Why another organisation: ncu of the generic kernel on this shape (profiles/r02_ncu_1_1_kernels.json) shows the load/store
// data pipe 88 % busy and the FP64 pipe 49 %: 2423 wavefronts per ciphertext and iteration, of which 517 are the 64 KB Fourier
// GGSW that EVERY one-ciphertext CTA reads for itself, 640 the exchange of the four transformed polynomials D_r between the
// teams, 512 the accumulator.  Here
//   * a team is 16 threads (R = 16), and a warp carries the SAME polynomial row r of TWO ciphertexts in its two half-warps:
//     both halves read the same key values at the same time, so one 256-byte request serves two ciphertexts (L2 traffic per
//     ciphertext halved; the data pipe still spends a wavefront per quarter-warp, see the profile);
//   * a team applies its own key row to its own D_r while that is still in registers (three foreign polynomials instead of four
//     are read back from shared memory);
//   * the accumulator coefficients a thread has just written stay in registers for the decomposition of the next iteration
//     (one shared-memory read of them less per iteration);
//   * warp = (pair, row): the three warps of a scheduler belong to three different ciphertext pairs that only meet at their own
//     128-thread named barriers, so their load and FP64 phases interleave.
// Shared memory per ciphertext: 4 x 4.25 KB exchange/Fourier regions + 16 KB accumulator + rotation amounts = 34.3 KB; six fit.
// Same mathematics, key layout and stage functions (br_core.cuh) as blind_rotate.cu; reference: fft64/crypto/bootstrap.rs:243-412,
// fft64/crypto/ggsw.rs:477-598.
#include "br_core.cuh"
#include "tmem_util.cuh"
#include "kernels.h"

namespace b200tfhe {

namespace {

constexpr int PR_CTS = 6, PR_R = 16, PR_K = 3, PR_K1 = 4, PR_L = 1;
constexpr int PR_THREADS = PR_CTS * PR_K1 * PR_R;      // 384 = 12 warps: warp w = (pair w >> 2, row w & 3)

template <typename T, int BL> struct PairConfig {
    typedef BRCore<T, PR_R, PR_K, PR_L, BL> C;
    static constexpr size_t FBUF_BYTES = (size_t)PR_K1 * C::REGION * sizeof(cplx);    // per ciphertext
    static constexpr size_t ACC_BYTES = (size_t)PR_K1 * C::N * sizeof(T);
    static constexpr size_t TW_BYTES = sizeof(TwTables<PR_R>);
    static size_t ahat_bytes(int n) { return (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15; }
    static size_t smem_bytes(int n) { return PR_CTS * (FBUF_BYTES + ACC_BYTES + ahat_bytes(n)) + TW_BYTES; }
};

template <typename T, int BL>
__global__ void __launch_bounds__(PR_THREADS, 1)
blind_rotate_pair_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                         const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                         size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                         const cplx *__restrict__ tw_inv_, int full_ctas, int tail_cts) {
#if defined(__CUDA_ARCH__)
    typedef PairConfig<T, BL> CFG;
    typedef typename CFG::C C;
    typedef TeamFFT<PR_R> FFT;
    constexpr int R = PR_R, K = PR_K, K1 = PR_K1, N = C::N, M = C::M, REGION = C::REGION;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & (R - 1), half = (threadIdx.x >> 4) & 1;
    const int r = warp & 3, pair = warp >> 2, ct_local = 2 * pair + half;
    // CTAs below full_ctas carry six ciphertexts; the rest -- the partial last round -- carry tail_cts (2 or 4) each, so that the
    // remainder is spread over all SMs: a CTA with one pair of ciphertexts takes well under half the time of a full one
    const int b = (int)blockIdx.x, cts_here = b < full_ctas ? PR_CTS : tail_cts;
    int ct = (b < full_ctas ? b * PR_CTS : full_ctas * PR_CTS + (b - full_ctas) * tail_cts) + ct_local;
    const bool pair_used = 2 * pair < cts_here && ct - half < batch;      // the pair's first ciphertext exists
    const bool live = ct_local < cts_here && ct < batch;
    if (!live) ct = batch - 1;      // ghost ciphertext (odd batch): same work, no output (keeps the pair's barriers uniform)

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    cplx *fbuf_ct = reinterpret_cast<cplx *>(smem_raw) + (size_t)ct_local * K1 * REGION;
    cplx *fbuf_r = fbuf_ct + (size_t)r * REGION;
    T *acc_r = reinterpret_cast<T *>(smem_raw + PR_CTS * CFG::FBUF_BYTES) + (size_t)(ct_local * K1 + r) * N;
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + PR_CTS * (CFG::FBUF_BYTES + CFG::ACC_BYTES) + ct_local * ahat_bytes);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + PR_CTS * (CFG::FBUF_BYTES + CFG::ACC_BYTES + ahat_bytes));
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += PR_THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    // ---- prologue: modulus switch of the whole input LWE (fft_impl/common.rs:26-43), acc <- LUT * X^{-b_hat} ---------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = r * R + lane; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    __syncthreads();
    if (!pair_used) return;         // an unused pair slot of a tail CTA: its four warps leave (the pairs only meet at their own barriers)
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = lane; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncwarp();
    const int bar_pair = 1 + pair;     // named barrier of this pair of ciphertexts: its four warps

    T own_re[R], own_im[R];            // the coefficients this thread rounds into: lane + R m, and + M
    static_for<0, R>([&](auto m) {
        own_re[decltype(m)::value] = acc_r[lane + R * decltype(m)::value];
        own_im[decltype(m)::value] = acc_r[lane + R * decltype(m)::value + M];
    });

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
        cplx x[R];
        // rotate-subtract + one-level signed decomposition (polynomial_algorithms.rs:425-490, decomposition.rs:27-86)
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
            const int c = lane + R * m;
            const T d0 = rot_read<T, N>(acc_r, c, a) - own_re[m];
            const T d1 = rot_read<T, N>(acc_r, c + M, a) - own_im[m];
            x[decltype(p)::value] = cplx{int_to_double(C::single_level_digit(d0)), int_to_double(C::single_level_digit(d1))};
        });
        FFT::fwd_pass1(lane, x, fbuf_r, twf);
        __syncwarp();
        FFT::fwd_pass2_load(lane, x, fbuf_r);
        // own key row (r, j = r): in flight during the second pass; register pp <-> Fourier position brev(pp)
        cplx F[R];
        {
            const cplx *row = ggsw_i + ((size_t)r * K1 + r) * M;
            static_for<0, R>([&](auto pp) { F[decltype(pp)::value] = ldg_stream(row + brev(decltype(pp)::value, FFT::LOGR) * R + lane); });
        }
        __syncwarp();
        FFT::fwd_pass2_compute(x);                  // x[p] = D_r at Fourier position p*R + lane
        static_for<0, R>([&](auto p) { fbuf_r[decltype(p)::value * R + lane] = x[decltype(p)::value]; });
        static_for<0, R>([&](auto pp) { F[decltype(pp)::value] = cmul(F[decltype(pp)::value], x[brev(decltype(pp)::value, FFT::LOGR)]); });
        // the three foreign rows, ascending; the key values of the next one are requested before the barrier / the arithmetic
        cplx h[R];
        const int r1 = (r == 0) ? 1 : 0, r2 = (r <= 1) ? 2 : 1, r3 = (r == 3) ? 2 : 3;
        auto load_row = [&](int rr) {
            const cplx *row = ggsw_i + ((size_t)rr * K1 + r) * M;
            static_for<0, R>([&](auto pp) { h[decltype(pp)::value] = ldg_stream(row + brev(decltype(pp)::value, FFT::LOGR) * R + lane); });
        };
        auto apply_row = [&](int rr) {
            const cplx *d = fbuf_ct + (size_t)rr * REGION;
            static_for<0, R>([&](auto pp) { C::mac_one(F[decltype(pp)::value], h[decltype(pp)::value], d[brev(decltype(pp)::value, FFT::LOGR) * R + lane]); });
        };
        load_row(r1);
        named_bar_sync(bar_pair, 128);              // every D of both ciphertexts of the pair is in shared memory
        apply_row(r1);
        load_row(r2);
        apply_row(r2);
        load_row(r3);
        apply_row(r3);
        FFT::inv_pass1_compute(lane, F, twi);       // butterflies and middle twiddle BEFORE the barrier: only the store needs the region free
        named_bar_sync(bar_pair, 128);              // every team has read every D: the regions are free for the inverse transform
        FFT::pass1_store(lane, F, fbuf_r);
        __syncwarp();
        FFT::inv_pass2(lane, x, fbuf_r);
        static_for<0, R>([&](auto m) {
            const int c = lane + R * decltype(m)::value;
            own_re[decltype(m)::value] = acc_r[c] + torus_from_double<T>(x[decltype(m)::value].re);
            own_im[decltype(m)::value] = acc_r[c + M] + torus_from_double<T>(x[decltype(m)::value].im);
            acc_r[c] = own_re[decltype(m)::value];
            acc_r[c + M] = own_im[decltype(m)::value];
        });
        __syncwarp();
    }

    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = lane; c < N; c += R) o[(size_t)r * N + c] = acc_r[c];
    } else {
        if (r < K) {
            for (int t = lane; t < N; t += R) o[(size_t)r * N + t] = (t == 0) ? acc_r[0] : (T)(0 - acc_r[N - t]);
        } else if (lane == 0) {
            o[(size_t)K * N] = acc_r[0];
        }
    }
#endif
}

}  // namespace

int br_pair_supported(const Params &p) {
    return p.bits == 64 && p.N == 512 && p.k == PR_K && p.pbs_level == 1 && p.pbs_base_log == 18;
}

int br_pair_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (!br_pair_supported(p)) return -1;
    if (a.batch <= 0) return 0;
    typedef PairConfig<uint64_t, 18> CFG;
    auto kern = blind_rotate_pair_kernel<uint64_t, 18>;
    const size_t smem = CFG::smem_bytes(a.n);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    // full rounds of six ciphertexts per SM, then the remainder in CTAs of one or two pairs over all SMs (a partial round of full CTAs
    // costs a whole round: 8 % of a Trivium wave of 8192)
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int full_ctas = a.batch / (PR_CTS * sms) * sms, rem = a.batch - full_ctas * PR_CTS;
    int tail_cts = 2 * ((rem + 2 * sms - 1) / (2 * sms));
    if (tail_cts < 2) tail_cts = 2;
    const int grid = full_ctas + (rem + tail_cts - 1) / tail_cts;
    kern<<<grid, PR_THREADS, smem, st>>>((const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx, (const cplx *)a.bsk,
                                         (uint64_t *)a.out, a.out_stride, a.n, a.batch, a.mode, (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv,
                                         full_ctas, tail_cts);
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

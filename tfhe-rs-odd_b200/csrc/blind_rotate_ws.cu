// blind_rotate_ws.cu -- blind rotation for (N = 2048, k = 1, l = 1), WARP-SPECIALISED: per ciphertext one "Fourier" warp and
// one "torus" warp that hand polynomials to each other through tensor memory.
//
// Why (profiles/r01_ncu_blind_rotate_summary.json, v8_tmem): in blind_rotate_tmem.cu the two warps of a ciphertext own one
// polynomial each and therefore run the SAME phase at the same time; they also share a scheduler (warps q and q+4 must
// share a TMEM lane quadrant, hence an SM sub-partition).  So the FP64 pipe of a sub-partition idles whenever its two warps
// decompose / accumulate / transpose, and the load/store pipe idles whenever they run FFT butterflies: FP64 pipe 59 % busy.
// Here the split is by KIND of work instead:
//   * Fourier warp F:  for r in {0,1}: X_r (TMEM) -> twist, radix-32 pass, transpose, radix-32 pass -> D_r (TMEM);
//                      for j in {0,1}: Y = sum_r G[r][j] * D_r (key chunks double-buffered in registers, L1-prefetched),
//                      inverse passes -> y_j (TMEM).  Pure FP64 stream, 4 transforms per iteration.
//   * torus warp T:    for r in {0,1}: y_r (TMEM) -> round to the torus, add into acc_r (shared memory), rotate by the next
//                      a_i, signed decomposition -> X_r (TMEM).  Pure load/store + integer stream.
// Both warps of a pair are busy with different pipes at the same time, and the hand-off is lane-local (thread l of both
// warps owns coefficients l + 32 m), so it costs no shared-memory bandwidth: tcgen05.st by the producer, mbarrier, tcgen05.ld
// by the consumer.  TMEM per ciphertext (one lane quadrant, 512 columns): D_0 | D_1 | y_0 / X_0 | y_1 / X_1.
// The arithmetic is the one of br_core.cuh / blind_rotate_tmem.cu, operation for operation.
//
// STATUS: experimental, off by default (B200TFHE_WS=1 selects it).  Parity-green, but 65.3 k PBS/s against 80.4 k for
// blind_rotate_tmem.cu (B200, batch 16384).  The phase trace (-DB200TFHE_WS_TRACE, tools/ws_trace.py) shows why: the
// Fourier warp is alone on the critical path (4 transforms x 3.0 k cycles + 2 MACs x 2.4-3.2 k cycles per iteration, the
// torus warp idles 60 %), and its MAC is bound by tensor-memory READ bandwidth (64 B/clk/SM: 32 tcgen05.ld.x32 per
// ciphertext-iteration instead of 8), not by the key loads -- prefetching the whole key block into a 60 KB L1
// (B200TFHE_WS_HALF=1 B200TFHE_WS_PF=1) does not help, and the half-size transpose it needs costs 8 %.  Two instruction
// streams per sub-partition also halve the instruction-cache hit rate when the Fourier warps are staggered (ring: 40-50 k).
#include "tmem_util.cuh"
#include "kernels.h"
#include <cstdio>
#include <cstdlib>

namespace b200tfhe {

constexpr int WS_CTS = 4;       // ciphertexts per CTA (one per TMEM lane quadrant)
constexpr int WS_COLS = 512;

#ifndef B200TFHE_WS_HALF
#define B200TFHE_WS_HALF 0      // 1: half-size transpose region (CTA under 196 KB, 60 KB of L1); 0: full-size region
#endif
#ifndef B200TFHE_WS_TWIST
#define B200TFHE_WS_TWIST 0     // 1: twist / un-twist applied by the torus warp; 0: by the Fourier warp
#endif
#ifndef B200TFHE_WS_PF
#define B200TFHE_WS_PF 0        // 1: whole 32 KB key block of a MAC prefetched one transform ahead; 0: rolling, 3 quarters ahead
#endif
#ifndef B200TFHE_WS_STAGGER
#define B200TFHE_WS_STAGGER 0   // 0: Fourier warps free-running (they stay within ~200 cycles of each other); 1/2: ring; -1: lock-step
#endif

#ifdef B200TFHE_WS_TRACE
constexpr int WTR_IT = 4, WTR_I0 = 300, WTR_ST = 12;
__device__ unsigned long long g_ws_trace[8][WTR_IT][WTR_ST];
#define WSTAMP(k) \
    if (blockIdx.x == 0 && lane == 0 && i >= WTR_I0 && i < WTR_I0 + WTR_IT) { \
        unsigned long long t_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_)::"memory"); g_ws_trace[warp][i - WTR_I0][k] = t_; }
#else
#define WSTAMP(k)
#endif

template <typename T, int BL>
__global__ void __launch_bounds__(WS_CTS * 64, 1)
blind_rotate_ws_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                       const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                       size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                       const cplx *__restrict__ tw_inv_) {
#if defined(__CUDA_ARCH__)
    constexpr int R = 32, K = 1, L = 1, K1 = 2;
    typedef BRCore<T, R, K, L, BL> C;
    typedef TeamFFT<R> FFT;
    constexpr int N = C::N, M = C::M, THREADS = WS_CTS * 64, Q8 = 8;      // Q8: positions per key quarter-chunk
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t tmem_base_smem;
    __shared__ __align__(8) uint64_t mbar[WS_CTS][4];      // per ciphertext: X_0 ready, X_1 ready, y_0 ready, y_1 ready

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3;                 // ciphertext slot = TMEM lane quadrant
    const bool fourier = warp < WS_CTS;     // warps 0..3: Fourier role; warps 4..7: torus role
    int ct = blockIdx.x * WS_CTS + q;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    constexpr int HREGION = B200TFHE_WS_HALF ? R * (R / 2 + 1) : FFT::REGION;   // half size keeps the CTA under 196 KB: 60 KB of L1
    constexpr size_t XCH_BYTES = (size_t)HREGION * sizeof(cplx), ACC_BYTES = (size_t)K1 * N * sizeof(T);
    cplx *region = reinterpret_cast<cplx *>(smem_raw + (size_t)q * XCH_BYTES);          // Fourier warp's transpose buffer
    auto transpose = [&](cplx (&x)[R]) {
#if B200TFHE_WS_HALF
        transpose_split<R>(lane, x, region);
#else
        FFT::pass1_store(lane, x, region);
        __syncwarp();
        FFT::fwd_pass2_load(lane, x, region);
        __syncwarp();
#endif
    };
    T *acc = reinterpret_cast<T *>(smem_raw + WS_CTS * XCH_BYTES + (size_t)q * ACC_BYTES);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + WS_CTS * (XCH_BYTES + ACC_BYTES) + q * ahat_bytes);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + WS_CTS * (XCH_BYTES + ACC_BYTES + ahat_bytes));
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr_u32(&tmem_base_smem)), "n"(WS_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x < WS_CTS * 4) mbarrier_init(smem_addr_u32(&mbar[threadIdx.x >> 2][threadIdx.x & 3]), 1);
    // ---- prologue: modulus switch, acc <- LUT * X^(-b^) (polynomial `warp >> 2` by each warp of the pair) ------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = (warp >> 2) * R + lane; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    tmem_fence_before();
    __syncthreads();
    tmem_fence_after();
    const uint32_t t_quad = tmem_base_smem + ((uint32_t)(q * 32) << 16);
    {
        const int r0 = warp >> 2;
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r0 * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = lane; c < N; c += R) acc[r0 * N + c] = rot_read<T, N>(lut, c, a0);
    }
    __syncthreads();
    const uint32_t bx0 = smem_addr_u32(&mbar[q][0]), by0 = smem_addr_u32(&mbar[q][2]);
    // TMEM slots of this ciphertext (column offsets): D_r at r*128, y_r / X_r at (2+r)*128

    if (fourier) {
        // =========================================== Fourier warp ====================================================
#if B200TFHE_WS_STAGGER > 0
        const int bar_top = 1 + q, bar_next = 1 + ((q + 1) & (WS_CTS - 1));     // 64 threads: 32 sync + 32 arrive
        if (q == WS_CTS - 1) named_bar_arrive(bar_next, 64);
#endif
        for (int i = 0; i < n; i++) {
            const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
            const uint32_t par = (uint32_t)(i & 1);
            WSTAMP(0)
#if B200TFHE_WS_STAGGER > 0
            named_bar_sync(bar_top, 64);
#elif B200TFHE_WS_STAGGER < 0
            named_bar_sync(1, 128);      // the 4 Fourier warps in lock-step
#endif
            // ---- forward transforms: X_r -> D_r ---------------------------------------------------------------------
#pragma unroll 1
            for (int r = 0; r < K1; r++) {
                cplx F[R];
                WSTAMP(1 + 3 * r)
                mbarrier_wait(bx0 + 8u * r, par);
                tmem_fence_after();
                WSTAMP(2 + 3 * r)
                static_for<0, 4>([&](auto part) {
                    uint32_t v[32];
                    tmem_ld32(t_quad + (2 + r) * 128 + decltype(part)::value * 32, v);
                    tmem_wait_ld();
                    static_for<0, 8>([&](auto e) { F[decltype(part)::value * 8 + decltype(e)::value] = cplx_from_words(v, decltype(e)::value); });
                });
                WSTAMP(3 + 3 * r)
#if B200TFHE_WS_STAGGER == 2
                if (r == 0 && (q != WS_CTS - 1 || i != n - 1)) named_bar_arrive(bar_next, 64);
#endif
                if (r == K1 - 1) {      // key block of the first MAC -> L1, one transform ahead (whole 32 KB, or its first 3 quarters)
                    static_for<0, (B200TFHE_WS_PF ? 8 : 3)>([&](auto s) { C::template quarter_prefetch<decltype(s)::value & 3>(lane, decltype(s)::value >> 2, 0, ggsw_i); });
                }
#if B200TFHE_WS_TWIST
                FFT::fwd_pass1_compute_twisted(lane, F, twf);      // the torus warp already applied the twist
#else
                FFT::fwd_pass1_compute(lane, F, twf);
#endif
#if B200TFHE_WS_STAGGER == 1
                if (r == 0 && (q != WS_CTS - 1 || i != n - 1)) named_bar_arrive(bar_next, 64);
#endif
                transpose(F);
                FFT::fwd_pass2_compute(F);
                // D_r -> TMEM in register order pp (position brev(pp)): the MAC's key chunks are then contiguous rows
                static_for<0, 4>([&](auto part) {
                    constexpr int b = decltype(part)::value * 8;
                    tmem_store_cplx8(t_quad + r * 128 + decltype(part)::value * 32, F[brev(b + 0, FFT::LOGR)], F[brev(b + 1, FFT::LOGR)],
                                     F[brev(b + 2, FFT::LOGR)], F[brev(b + 3, FFT::LOGR)], F[brev(b + 4, FFT::LOGR)],
                                     F[brev(b + 5, FFT::LOGR)], F[brev(b + 6, FFT::LOGR)], F[brev(b + 7, FFT::LOGR)]);
                });
            }
            tmem_wait_st();
            WSTAMP(7)
            // ---- Y_j = sum_r G[r][j] * D_r, inverse transform, y_j -> TMEM ---------------------------------------------
#pragma unroll 1
            for (int j = 0; j < K1; j++) {
                cplx Y[R];
                cplx h[2][Q8];
                // 8 steps: s = r*4 + quarter; the key quarter of step s+1 is loaded while step s is consumed
                C::template quarter_load<0>(lane, 0, j, ggsw_i, h[0]);
                static_for<0, 2 * 4>([&](auto ss) {
                    constexpr int s = decltype(ss)::value, r = s >> 2, qt = s & 3;
                    if constexpr (s + 1 < 8) C::template quarter_load<(s + 1) & 3>(lane, (s + 1) >> 2, j, ggsw_i, h[(s + 1) & 1]);
#if !B200TFHE_WS_PF
                    if constexpr (s + 3 < 8) C::template quarter_prefetch<(s + 3) & 3>(lane, (s + 3) >> 2, j, ggsw_i);
#endif
                    uint32_t v[32];
                    tmem_ld32(t_quad + r * 128 + qt * 32, v);
                    tmem_wait_ld();
                    static_for<0, Q8>([&](auto e) {
                        constexpr int pp = qt * Q8 + decltype(e)::value;
                        const cplx d = cplx_from_words(v, decltype(e)::value);
                        if constexpr (r == 0) Y[pp] = cmul(d, h[s & 1][decltype(e)::value]);
                        else C::mac_one(Y[pp], h[s & 1][decltype(e)::value], d);
                    });
                });
                if (j + 1 < K1) {       // key block of the next MAC -> L1 during this inverse transform
                    static_for<0, (B200TFHE_WS_PF ? 8 : 3)>([&](auto s) { C::template quarter_prefetch<decltype(s)::value & 3>(lane, decltype(s)::value >> 2, j + 1, ggsw_i); });
                }
                asm volatile("" : "+d"(Y[0].re), "+d"(Y[31].im));
                WSTAMP(8 + 2 * j)
                FFT::inv_pass1_compute(lane, Y, twi);
                transpose(Y);
                fft_dit<R, +1>(Y);
#if !B200TFHE_WS_TWIST
                static_for<0, R>([&](auto m) { Y[decltype(m)::value] = cmul_root<-decltype(m)::value * FFT::TWQ>(Y[decltype(m)::value]); });
#endif
                static_for<0, 4>([&](auto part) {
                    constexpr int b = decltype(part)::value * 8;
                    tmem_store_cplx8(t_quad + (2 + j) * 128 + decltype(part)::value * 32, Y[b + 0], Y[b + 1], Y[b + 2], Y[b + 3],
                                     Y[b + 4], Y[b + 5], Y[b + 6], Y[b + 7]);
                });
                tmem_wait_st();
                tmem_fence_before();
                __syncwarp();
                if (lane == 0) mbarrier_arrive(by0 + 8u * j);
                WSTAMP(9 + 2 * j)
            }
        }
    } else {
        // ============================================ torus warp =====================================================
        const unsigned char *acc_bytes = reinterpret_cast<const unsigned char *>(acc);
        // decomposition of acc_r * X^a - acc_r at this lane's 32 folded points -> X_r (TMEM), then signal the Fourier warp
        auto decompose = [&](int r, int i, const T (&own_re)[R], const T (&own_im)[R]) {
            const uint32_t base8 = ((uint32_t)(lane - (int)ahat[i]) << 3);
            const unsigned char *poly = acc_bytes + (size_t)r * N * sizeof(T);
            static_for<0, 4>([&](auto part) {
                cplx x[8];
                static_for<0, 8>([&](auto e) {
                    constexpr int p = decltype(part)::value * 8 + decltype(e)::value, m = brev(p, FFT::LOGR);
                    x[decltype(e)::value] = cplx{rot_digit_f64<BL>(poly, base8 + 8u * R * m, own_re[m]),
                                                 rot_digit_f64<BL>(poly, base8 + 8u * R * m + 8u * M, own_im[m])};
                });
#if B200TFHE_WS_TWIST
                static_for<0, 8>([&](auto e) {      // forward twist e^{i pi m / (2R)} (compile-time root), as TeamFFT::twist
                    constexpr int m = brev(decltype(part)::value * 8 + decltype(e)::value, FFT::LOGR);
                    x[decltype(e)::value] = cmul_root<m * FFT::TWQ>(x[decltype(e)::value]);
                });
#endif
                tmem_store_cplx8(t_quad + (2 + r) * 128 + decltype(part)::value * 32, x[0], x[1], x[2], x[3], x[4], x[5], x[6], x[7]);
            });
            tmem_wait_st();
            tmem_fence_before();
            __syncwarp();
            if (lane == 0) mbarrier_arrive(bx0 + 8u * r);
        };
        // iteration 0: the accumulators come straight from the prologue
#pragma unroll 1
        for (int r = 0; r < K1; r++) {
            T own_re[R], own_im[R];
            static_for<0, R>([&](auto m) {
                own_re[decltype(m)::value] = acc[r * N + lane + R * decltype(m)::value];
                own_im[decltype(m)::value] = acc[r * N + lane + R * decltype(m)::value + M];
            });
            decompose(r, 0, own_re, own_im);
        }
        for (int i = 0; i < n; i++) {
            const uint32_t par = (uint32_t)(i & 1);
#pragma unroll 1
            for (int r = 0; r < K1; r++) {
                T own_re[R], own_im[R];
                T *acc_r = acc + r * N;
                WSTAMP(0 + 4 * r)
                mbarrier_wait(by0 + 8u * r, par);
                tmem_fence_after();
                WSTAMP(1 + 4 * r)
                static_for<0, 4>([&](auto part) {
                    uint32_t v[32];
                    tmem_ld32(t_quad + (2 + r) * 128 + decltype(part)::value * 32, v);
                    tmem_wait_ld();
                    static_for<0, 8>([&](auto e) {
                        constexpr int m = decltype(part)::value * 8 + decltype(e)::value;
#if B200TFHE_WS_TWIST
                        const cplx t = cmul_root<-m * FFT::TWQ>(cplx_from_words(v, decltype(e)::value));      // un-twist
#else
                        const cplx t = cplx_from_words(v, decltype(e)::value);
#endif
                        const int c = lane + R * m;
                        own_re[m] = acc_r[c] + torus_from_double<T>(t.re);
                        own_im[m] = acc_r[c + M] + torus_from_double<T>(t.im);
                        acc_r[c] = own_re[m];
                        acc_r[c + M] = own_im[m];
                    });
                });
                __syncwarp();
                WSTAMP(2 + 4 * r)
                if (i + 1 < n) decompose(r, i + 1, own_re, own_im);
                WSTAMP(3 + 4 * r)
            }
        }
    }

    tmem_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(WS_COLS));
    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    const int r = warp >> 2;      // each warp of the pair writes one polynomial
    const T *acc_r = acc + r * N;
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

int br_ws_supported(const Params &p) {
    return p.bits == 64 && p.N == 2048 && p.k == 1 && p.pbs_level == 1 && p.pbs_base_log == 23;
}

int br_ws_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (!br_ws_supported(p)) return -1;
    typedef BRCore<uint64_t, 32, 1, 1, 23> C;
    auto kern = blind_rotate_ws_kernel<uint64_t, 23>;
    size_t ahat = (((size_t)(a.n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    size_t smem = WS_CTS * ((size_t)(B200TFHE_WS_HALF ? 32 * 17 : C::REGION) * sizeof(cplx) + (size_t)2 * C::N * sizeof(uint64_t) + ahat) + sizeof(TwTables<32>);
    // with the half-size transpose region the CTA stays under 196 KB; the carve-out hint leaves the rest of the array to L1
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem, (int)((smem + 2048) * 100 / (228 * 1024)) + 1)) return (int)e;
    int grid = (a.batch + WS_CTS - 1) / WS_CTS;
    kern<<<grid, WS_CTS * 64, smem, st>>>((const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx,
                                         (const cplx *)a.bsk, (uint64_t *)a.out, a.out_stride, a.n, a.batch, a.mode,
                                         (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv);
#ifdef B200TFHE_WS_TRACE
    if (const char *f = getenv("B200TFHE_TRACE_FILE")) {
        static unsigned long long host[8][WTR_IT][WTR_ST];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(host, g_ws_trace, sizeof(host));
        if (FILE *fp = fopen(f, "w")) {
            for (int w = 0; w < 8; w++)
                for (int it = 0; it < WTR_IT; it++) {
                    fprintf(fp, "%d %d", w, it);
                    for (int k = 0; k < WTR_ST; k++) fprintf(fp, " %llu", host[w][it][k]);
                    fprintf(fp, "\n");
                }
            fclose(fp);
        }
    }
#endif
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

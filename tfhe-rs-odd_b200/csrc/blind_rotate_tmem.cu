// blind_rotate_tmem.cu -- blind rotation for (N = 2048, k = 1, l = 1) with the forward transforms exchanged through
// TENSOR MEMORY: 4 ciphertexts per CTA (8 warps / SM) instead of 3 (6 warps).
//
// Why: the shared-memory budget of blind_rotate.cu is 67 KB per ciphertext (accumulator 32 KB + one exchange/Fourier region
// per polynomial 2 x 17 KB), i.e. 3 ciphertexts = 6 warps per SM, and ncu shows the FP64 pipe starved by latency at that
// occupancy (profiles/r01_ncu_blind_rotate_summary.json).  Here
//   * a team's own Fourier polynomial D_r never leaves its registers: its own key row is applied register-to-register;
//   * D_r reaches the other team through TMEM (tcgen05.st / tcgen05.ld, 32x32b shape: thread l of a warp owns TMEM lane
//     32*(warp%4)+l).  The two teams of ciphertext q are warps q and q+4, which address the same TMEM lane quadrant, and the
//     MAC needs exactly "same lane, same Fourier position" of the other polynomial -- no cross-lane movement at all;
//   * every team owns a HALF-size (8.5 KB) transpose region and does each 32x32 transpose in two column halves (16 source
//     lanes at a time, 16 values parked in registers in between), so a ciphertext needs 32 + 17 = 49 KB of shared memory,
//     4 fit next to the twiddle tables, and the teams never wait for each other's transposes;
//   * the TMEM slots are double-buffered by iteration parity, so ONE 64-thread named barrier per iteration and ciphertext
//     is the only inter-warp synchronisation of the data path.
// TMEM use per CTA: 4 quadrants x 2 buffers x 2 polynomials x 128 columns = all 512 columns (1 CTA per SM).  No
// tensor-core instruction is involved: TMEM is a 256 KB per-SM scratchpad that costs no shared-memory capacity or bandwidth.
#include "tmem_util.cuh"
#include "kernels.h"
#include <cstdio>
#include <cstdlib>

namespace b200tfhe {

// Iteration alignment of the 4 ciphertexts of a CTA:
//   0  none (warps drift apart: 4 instruction streams, instruction-cache misses)
//   -1 lockstep (__syncthreads per iteration: one instruction stream, but every ciphertext hits the shared-memory pipe in
//      the same cycles and the FP64 pipes in the same cycles -- measured: time = T_lsu + T_fp64)
//   k>0 STAGGERED ring: ciphertext q starts iteration i when ciphertext q-1 has decomposed k of its 32 points of iteration
//      i (and 0 waits for 3 of iteration i-1).  The streams stay a few hundred instructions apart (inside the 32 KB L1.5
//      instruction cache) while the load/store bursts of one ciphertext overlap the FP64 phases of the others.
#ifndef B200TFHE_TMEM_FENCE
#define B200TFHE_TMEM_FENCE 4
#endif
#ifndef B200TFHE_TMEM_PREFETCH
#define B200TFHE_TMEM_PREFETCH 2
#endif
#ifndef B200TFHE_TMEM_STAGGER
#define B200TFHE_TMEM_STAGGER 8
#endif

// Debug build only (-DB200TFHE_TMEM_TRACE): CTA 0 records %clock64 at the phase boundaries of a few iterations; the
// launcher dumps them to the file named by B200TFHE_TRACE_FILE (see tools/tmem_trace.py).
#ifdef B200TFHE_TMEM_TRACE
constexpr int TR_IT = 6, TR_I0 = 300, TR_ST = 12;
__device__ unsigned long long g_tmem_trace[8][TR_IT][TR_ST];
#define STAMP(k) \
    if (blockIdx.x == 0 && lane == 0 && i >= TR_I0 && i < TR_I0 + TR_IT) { \
        unsigned long long t_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_)::"memory"); g_tmem_trace[warp][i - TR_I0][k] = t_; }
#else
#define STAMP(k)
#endif

constexpr int TM_CTS = 4;        // ciphertexts per CTA
constexpr int TM_COLS = 512;     // TMEM columns allocated per CTA (2 buffers x 2 polynomials x 128)

template <typename T, int BL>
__global__ void __launch_bounds__(TM_CTS * 64, 1)
blind_rotate_tmem_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                         const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                         size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                         const cplx *__restrict__ tw_inv_) {
#if defined(__CUDA_ARCH__)
    constexpr int R = 32, K = 1, L = 1, K1 = 2;
    typedef BRCore<T, R, K, L, BL> C;
    typedef TeamFFT<R> FFT;
    constexpr int N = C::N, CH = C::CH, THREADS = TM_CTS * 64;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, r = warp >> 2;        // ciphertext slot = TMEM lane quadrant; team (polynomial) index
    int ct = blockIdx.x * TM_CTS + q;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    constexpr int HREGION = R * (R / 2 + 1);      // cplx per private half-size transpose region
    constexpr size_t XCH_BYTES = (size_t)HREGION * sizeof(cplx), ACC_BYTES = (size_t)N * sizeof(T);
    cplx *hreg = reinterpret_cast<cplx *>(smem_raw + (size_t)warp * XCH_BYTES);
    T *acc_r = reinterpret_cast<T *>(smem_raw + 2 * TM_CTS * XCH_BYTES + (size_t)(q * K1 + r) * ACC_BYTES);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + 2 * TM_CTS * (XCH_BYTES + ACC_BYTES) + q * ahat_bytes);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + 2 * TM_CTS * (XCH_BYTES + ACC_BYTES) + TM_CTS * ahat_bytes);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    // ---- tensor-memory allocation (one warp), address broadcast through shared memory -------------------------------
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "n"(TM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // ---- prologue ----------------------------------------------------------------------------------------------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = r * R + lane; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    tmem_fence_before();
    __syncthreads();
    tmem_fence_after();
    // lane field (bits 31..16) = first lane of this warp's quadrant; columns: [buffer][polynomial][128]
    const uint32_t t_quad = tmem_base_smem + ((uint32_t)(q * 32) << 16);
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = lane; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncwarp();
    const int bar_ct = 1 + q;      // named barrier of this ciphertext (64 threads)
#if B200TFHE_TMEM_STAGGER > 0
    const int bar_top = 1 + TM_CTS + q, bar_next = 1 + TM_CTS + ((q + 1) & (TM_CTS - 1));   // 128 threads: 64 sync + 64 arrive
    if (q == TM_CTS - 1) named_bar_arrive(bar_next, 128);     // lets ciphertext 0 into iteration 0
#endif

    // The thread's own 64 updated coefficients stay in registers from the accumulate step to the decomposition that follows
    // it (they are re-read from shared memory one iteration later; the other lanes read them rotated).
    T own_re[R], own_im[R];
    static_for<0, R>([&](auto m) {
        own_re[decltype(m)::value] = acc_r[lane + R * decltype(m)::value];
        own_im[decltype(m)::value] = acc_r[lane + R * decltype(m)::value + C::M];
    });
    const unsigned char *acc_bytes = reinterpret_cast<const unsigned char *>(acc_r);
    for (int i = 0; i < n; i++) {
        // decomposition of acc*X^a_i - acc at this lane's 32 folded points -> F (doubles), with the ring alignment inside
        cplx F[R];
        STAMP(0)
#if B200TFHE_TMEM_STAGGER < 0
        __syncthreads();
#elif B200TFHE_TMEM_STAGGER > 0
        named_bar_sync(bar_top, 128);
#endif
        STAMP(1)
        const uint32_t base8 = (uint32_t)(lane - (int)ahat[i]) << 3;      // byte offset of coefficient (lane - a); bit 14 = sign
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
#if B200TFHE_TMEM_STAGGER > 0
            if constexpr (decltype(p)::value == B200TFHE_TMEM_STAGGER) {
                if (q != TM_CTS - 1 || i != n - 1) named_bar_arrive(bar_next, 128);
            }
#endif
            if constexpr (decltype(p)::value % B200TFHE_TMEM_FENCE == 0) asm volatile("" ::: "memory");   // bounds the loads in flight (registers)
            F[decltype(p)::value] = cplx{rot_digit_f64<BL>(acc_bytes, base8 + 8u * R * m, own_re[m]),
                                         rot_digit_f64<BL>(acc_bytes, base8 + 8u * R * m + 8u * C::M, own_im[m])};
        });
        const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
        const uint32_t t_mine = t_quad + (uint32_t)((i & 1) * 256 + r * 128);
        const uint32_t t_other = t_quad + (uint32_t)((i & 1) * 256 + (1 - r) * 128);
        STAMP(2)
        FFT::fwd_pass1_compute(lane, F, twf);
        asm volatile("" : "+d"(F[0].re), "+d"(F[31].im));
        STAMP(3)
        transpose_split<R>(lane, F, hreg);
        STAMP(4)
        cplx g[CH];
        C::template own_chunk_load<0>(lane, r, ggsw_i, g);      // key loads in flight during the second radix-32 pass
#if B200TFHE_TMEM_PREFETCH == 5
        if (q == 0) { C::template chunk_prefetch<1>(lane, r, r, ggsw_i); C::template chunk_prefetch<0>(lane, 1 - r, r, ggsw_i); }
#elif B200TFHE_TMEM_PREFETCH >= 1
        C::template chunk_prefetch<1>(lane, r, r, ggsw_i);
#endif
#if B200TFHE_TMEM_PREFETCH >= 2 && B200TFHE_TMEM_PREFETCH != 5
        C::template chunk_prefetch<0>(lane, 1 - r, r, ggsw_i);
#endif
        FFT::fwd_pass2_compute(F);                              // F[p] = D_r at Fourier position p*R + lane
        // D_r -> tensor memory in register order pp (position brev(pp)), so that the reader's chunks are contiguous columns
        static_for<0, 4>([&](auto part) {
            uint32_t v[32];
            static_for<0, 8>([&](auto e) {
                constexpr int pos = brev(decltype(part)::value * 8 + decltype(e)::value, FFT::LOGR);
                v[4 * decltype(e)::value + 0] = (uint32_t)__double2loint(F[pos].re);
                v[4 * decltype(e)::value + 1] = (uint32_t)__double2hiint(F[pos].re);
                v[4 * decltype(e)::value + 2] = (uint32_t)__double2loint(F[pos].im);
                v[4 * decltype(e)::value + 3] = (uint32_t)__double2hiint(F[pos].im);
            });
            tmem_st32(t_mine + decltype(part)::value * 32, v);
        });
        // own key row, register to register and in place: Y[pp] = G[r][r][brev(pp)] * D_r[brev(pp)]
        cplx Y[R];
        {
            cplx xb[R];
            static_for<0, R>([&](auto pp) { xb[decltype(pp)::value] = F[brev(decltype(pp)::value, FFT::LOGR)]; });
            C::template own_chunk_apply<0>(xb, g, Y);
            C::template own_chunk_load<1>(lane, r, ggsw_i, g);
            C::template own_chunk_apply<1>(xb, g, Y);
        }
        cplx h[CH];
        C::template other_chunk_load<0>(lane, 1 - r, r, ggsw_i, h);   // in flight across the barrier
#if B200TFHE_TMEM_PREFETCH == 3
        C::template chunk_prefetch<1>(lane, 1 - r, r, ggsw_i);
#endif
        tmem_wait_st();
        tmem_fence_before();
        STAMP(5)
        named_bar_sync(bar_ct, 64);
        tmem_fence_after();
#if B200TFHE_TMEM_PREFETCH == 4
        C::template chunk_prefetch<1>(lane, 1 - r, r, ggsw_i);
#endif
        STAMP(6)
        static_for<0, 4>([&](auto part) {
            if constexpr (decltype(part)::value == 2) C::template other_chunk_load<1>(lane, 1 - r, r, ggsw_i, h);
            uint32_t v[32];
            tmem_ld32(t_other + decltype(part)::value * 32, v);
            tmem_wait_ld();
            static_for<0, 8>([&](auto e) {
                constexpr int pp = decltype(part)::value * 8 + decltype(e)::value;
                cplx d{__hiloint2double((int)v[4 * decltype(e)::value + 1], (int)v[4 * decltype(e)::value + 0]),
                       __hiloint2double((int)v[4 * decltype(e)::value + 3], (int)v[4 * decltype(e)::value + 2])};
                C::mac_one(Y[pp], h[pp % CH], d);
                asm volatile("" : "+d"(Y[pp].re), "+d"(Y[pp].im));   // consume this part before the next tcgen05.ld (registers)
            });
        });
        // no second barrier: the slots are double-buffered, and the next write of THIS buffer (iteration i+2) comes after
        // the barrier of iteration i+1, which the other team reaches only after these reads
        STAMP(7)
        FFT::inv_pass1_compute(lane, Y, twi);
        asm volatile("" : "+d"(Y[0].re), "+d"(Y[31].im));
        STAMP(8)
        transpose_split<R>(lane, Y, hreg);
        STAMP(9)
        fft_dit<R, +1>(Y);
        asm volatile("" : "+d"(Y[0].re), "+d"(Y[31].im));
        STAMP(10)
        static_for<0, R>([&](auto m) {
            cplx t = cmul_root<-decltype(m)::value * FFT::TWQ>(Y[decltype(m)::value]);
            const int c = lane + R * decltype(m)::value;
            if constexpr (decltype(m)::value % B200TFHE_TMEM_FENCE == 0) asm volatile("" ::: "memory");
            own_re[decltype(m)::value] = acc_r[c] + torus_from_double<T>(t.re);
            own_im[decltype(m)::value] = acc_r[c + C::M] + torus_from_double<T>(t.im);
            acc_r[c] = own_re[decltype(m)::value];
            acc_r[c + C::M] = own_im[decltype(m)::value];
        });
        __syncwarp();
        STAMP(11)
    }

    tmem_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(TM_COLS));
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

int br_tmem_supported(const Params &p) {
    return p.bits == 64 && p.N == 2048 && p.k == 1 && p.pbs_level == 1 && p.pbs_base_log == 23;
}

int br_tmem_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (!br_tmem_supported(p)) return -1;
    typedef BRCore<uint64_t, 32, 1, 1, 23> C;
    auto kern = blind_rotate_tmem_kernel<uint64_t, 23>;
    size_t ahat = (((size_t)(a.n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    size_t smem = TM_CTS * ((size_t)2 * 32 * 17 * sizeof(cplx) + (size_t)2 * C::N * sizeof(uint64_t) + ahat) + sizeof(TwTables<32>);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    int grid = (a.batch + TM_CTS - 1) / TM_CTS;
    kern<<<grid, TM_CTS * 64, smem, st>>>((const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx,
                                         (const cplx *)a.bsk, (uint64_t *)a.out, a.out_stride, a.n, a.batch, a.mode,
                                         (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv);
#ifdef B200TFHE_TMEM_TRACE
    if (const char *f = getenv("B200TFHE_TRACE_FILE")) {
        static unsigned long long host[8][TR_IT][TR_ST];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(host, g_tmem_trace, sizeof(host));
        if (FILE *fp = fopen(f, "w")) {
            for (int w = 0; w < 8; w++)
                for (int it = 0; it < TR_IT; it++) {
                    fprintf(fp, "%d %d", w, it);
                    for (int k = 0; k < TR_ST; k++) fprintf(fp, " %llu", host[w][it][k]);
                    fprintf(fp, "\n");
                }
            fclose(fp);
        }
    }
#endif
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

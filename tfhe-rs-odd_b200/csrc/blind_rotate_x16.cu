// blind_rotate_x16.cu -- blind rotation for the N = 512 sets with several decomposition levels (boolean DEFAULT_PARAMETERS:
// k = 2, l = 3, u32) with the Fourier polynomials exchanged through TENSOR MEMORY: 8 ciphertexts per CTA.
//
// Why: on the generic kernel (blind_rotate.cu) this shape is bound by the shared-memory pipe, not by FP64
// (profiles/r01_ncu_other_kernels.json: l1tex data pipe 87 %, FP64 35 %): every team re-reads ALL (k+1) l Fourier polynomials
// D[r][level] of its ciphertext from shared memory for the multiply-accumulate (108 KB of the 370 KB a ciphertext moves per
// iteration) after each team stored its own (36 KB).  Here
//   * a team is a HALF-warp (R = 16 threads x 16 points, M = 256 = 16 x 16); the k+1 teams of a ciphertext sit at the SAME lanes of
//     the warps q, q+4, q+8 (same tensor-memory lane quadrant), so "the other team's D at my lane" is a plain tcgen05.ld -- the
//     multiply-accumulate takes nothing from shared memory.  The two half-warps of a warp are two different ciphertexts;
//   * levels run one after the other (rolled loop): decompose level t, forward transform, publish D[r][t] (64 columns), ONE named
//     barrier over the (k+1) warps, accumulate F_r += sum_r' G[t][r'][r] (.) D[r'][t] with F in registers; D slots are double
//     buffered by the running level count, so there is no second barrier;
//   * both half-warps of a warp read the SAME bootstrap-key elements in the same instruction (lanes l and l + 16 coincide), so
//     the key's L1 traffic per ciphertext halves.
// Summation order is the reference's (level l down to 1, row ascending: fft64/crypto/ggsw.rs:524-577); arithmetic stages are the
// shared ones of br_core.cuh / fft_reg.cuh, so the lane emulator (tests/native/emul.cpp) describes this kernel too.
#include "tmem_util.cuh"
#include "kernels.h"

namespace b200tfhe {

constexpr int X16_CTS = 8;       // ciphertexts per CTA: 4 lane quadrants x 2 half-warps
constexpr int X16_COLS = 512;    // tensor-memory columns allocated (2 buffers x (k+1) teams x 64 used)

// VAR (selected by the context option "x16_variant"): bit 0 = the level's bootstrap-key block (all (k+1)^2 rows of one level: 36 KB for the
// boolean set, the SAME for the 8 ciphertexts of the CTA) is staged into shared memory by the TMA engine (cp.async.bulk + mbarrier) while
// the level's transforms run, and the multiply-accumulate reads it with LDS instead of waiting for L2; bit 1 = both halves of a D row are
// requested from tensor memory before the first is consumed.
// Measured and rejected on this kernel (B200, batch 8192, generic kernel 104.1 ms, this kernel 107.4 ms): an L1 prefetch of the level's key
// rows (119.8 ms: the 58 KB of L1 left beside 169 KB of shared memory thrash) and parking half of F in tensor memory during the transform
// (124.5 ms).
template <typename T, int K, int L, int BL, int VAR>
__global__ void __launch_bounds__(4 * (K + 1) * 32, 1)
blind_rotate_x16_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                        const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                        size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                        const cplx *__restrict__ tw_inv_) {
#if defined(__CUDA_ARCH__)
    constexpr int R = 16, K1 = K + 1, THREADS = 4 * K1 * 32;
    static_assert(2 * K1 * 64 <= X16_COLS, "D slots of the k+1 teams, double buffered, must fit 512 tensor-memory columns");
    typedef BRCore<T, R, K, L, BL> C;
    typedef TeamFFT<R> FFT;
    typedef typename StateWord<T, L * BL>::type S;
    constexpr int N = C::N, M = C::M, REGION = C::REGION;
    constexpr bool TMA_KEY = (VAR & 1) != 0, PAIR_LD = (VAR & 2) != 0;
    constexpr uint32_t LEVEL_BYTES = (uint32_t)(K1 * K1 * M * sizeof(cplx));      // one level of one Fourier GGSW
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, r = warp >> 2;            // tensor-memory lane quadrant; team (polynomial) index
    const int h = lane >> 4, l16 = lane & 15;         // half-warp = ciphertext of the pair; lane inside the team
    const int slot = 2 * q + h;
    int ct = blockIdx.x * X16_CTS + slot;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;                        // ghost ciphertext: same work, no output (keeps the barriers uniform)

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    constexpr size_t REG_BYTES = (size_t)REGION * sizeof(cplx), ACC_BYTES = (size_t)N * sizeof(T);
    cplx *region = reinterpret_cast<cplx *>(smem_raw + (size_t)(slot * K1 + r) * REG_BYTES);
    T *acc_r = reinterpret_cast<T *>(smem_raw + X16_CTS * K1 * REG_BYTES + (size_t)(slot * K1 + r) * ACC_BYTES);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + X16_CTS * K1 * (REG_BYTES + ACC_BYTES) + slot * ahat_bytes);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + X16_CTS * K1 * (REG_BYTES + ACC_BYTES) + X16_CTS * ahat_bytes);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    // TMA-staged key: [level block | mbarrier | counter of the warps that are done with the block] after the twiddle tables
    unsigned char *key_raw = reinterpret_cast<unsigned char *>(tw_s + 2 * TW_HALF);
    const cplx *key_s = reinterpret_cast<const cplx *>(key_raw);
    const uint32_t key_sa = smem_addr_u32(key_raw), full_bar = smem_addr_u32(key_raw + LEVEL_BYTES);
    unsigned int *done_cnt = reinterpret_cast<unsigned int *>(key_raw + LEVEL_BYTES + 8);
    const uint32_t total_steps = (uint32_t)n * L;
    auto fetch_level = [&](uint32_t s_) {          // one thread: block of running level s_ = (iteration s_/L, level index L-1 - s_%L)
        const unsigned char *src = reinterpret_cast<const unsigned char *>(bsk + (size_t)(s_ / L) * C::GGSW_CPLX + (size_t)(L - 1 - s_ % L) * K1 * K1 * M);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full_bar), "r"(LEVEL_BYTES) : "memory");
        static_for<0, K1>([&](auto c) {
            constexpr uint32_t part = LEVEL_BYTES / K1;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(key_sa + decltype(c)::value * part),
                         "l"(src + decltype(c)::value * part), "r"(part), "r"(full_bar)
                         : "memory");
        });
    };
    if constexpr (TMA_KEY) {
        if (threadIdx.x == 0) {
            mbarrier_init(full_bar, 1);
            *done_cnt = 0;
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "n"(X16_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // ---- prologue: modulus switch of the whole input (fft_impl/common.rs:26-43), acc <- LUT * X^{-b_hat} -----------------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = r * R + l16; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    tmem_fence_before();
    __syncthreads();
    tmem_fence_after();
    const uint32_t t_quad = tmem_base_smem + ((uint32_t)(q * 32) << 16);
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = l16; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncwarp();
    if constexpr (TMA_KEY) {
        if (threadIdx.x == 0 && total_steps > 0) fetch_level(0);      // after the __syncthreads above: barrier initialised, visible
    }
    const int bar_q = 1 + q;                          // named barrier of the K1 warps of this lane quadrant
    uint32_t step = 0;                                // running level count: D slot buffer = step & 1

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
        cplx F[R];                                    // F[pp] <-> Fourier position brev(pp) * R + l16 (what the inverse DIT wants)
        static_for<0, R>([&](auto pp) { F[decltype(pp)::value] = cplx{0.0, 0.0}; });
#pragma unroll 1
        for (int t = 0; t < L; t++, step++) {         // t = 0 <-> level L, the first digit the recurrence emits
            const cplx *glvl = ggsw_i + (size_t)(L - 1 - t) * K1 * K1 * M + (size_t)r * M + l16;
            {
                cplx x[R];
                static_for<0, R>([&](auto p) {
                    constexpr int m = brev(decltype(p)::value, FFT::LOGR);
                    const int c = l16 + R * m;
                    S s0 = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c, a) - acc_r[c], BL, L);
                    S s1 = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M], BL, L);
                    int32_t g0 = 0, g1 = 0;
                    static_for<0, L>([&](auto e) {    // every level's digit is a few integer instructions: compute all, select one
                        const int32_t e0 = (int32_t)decomp_next_digit<S>(s0, BL), e1 = (int32_t)decomp_next_digit<S>(s1, BL);
                        if (t == decltype(e)::value) { g0 = e0; g1 = e1; }
                    });
                    x[decltype(p)::value] = cplx{int_to_double(g0), int_to_double(g1)};
                });
                FFT::fwd_pass1(l16, x, region, twf);
                __syncwarp();
                FFT::fwd_pass2_load(l16, x, region);
                __syncwarp();
                FFT::fwd_pass2_compute(x);            // x[p] = D_r[t] at Fourier position p * R + l16
                const uint32_t t_mine = t_quad + (uint32_t)(((step & 1) * K1 + r) * 64);
                static_for<0, 2>([&](auto half) {
                    uint32_t v[32];
                    static_for<0, 8>([&](auto e) {
                        constexpr int pos = brev(decltype(half)::value * 8 + decltype(e)::value, FFT::LOGR);
                        v[4 * decltype(e)::value + 0] = (uint32_t)__double2loint(x[pos].re);
                        v[4 * decltype(e)::value + 1] = (uint32_t)__double2hiint(x[pos].re);
                        v[4 * decltype(e)::value + 2] = (uint32_t)__double2loint(x[pos].im);
                        v[4 * decltype(e)::value + 3] = (uint32_t)__double2hiint(x[pos].im);
                    });
                    tmem_st32(t_mine + decltype(half)::value * 32, v);
                });
            }
            tmem_wait_st();
            tmem_fence_before();
            named_bar_sync(bar_q, K1 * 32);
            tmem_fence_after();
            // F_r += sum_r' G[level][r'][r] (.) D[r'][t]: rows ascending (the reference's order); every D comes from tensor memory (the
            // team's own too: its registers are free for the key values)
            const uint32_t t_buf = t_quad + (uint32_t)((step & 1) * K1 * 64);
            if constexpr (TMA_KEY) mbarrier_wait(full_bar, step & 1);        // this level's key block has landed in shared memory
#pragma unroll
            for (int rr = 0; rr < K1; rr++) {
                const cplx *row = TMA_KEY ? key_s + (size_t)(rr * K1 + r) * M + l16 : glvl + (size_t)rr * K1 * M;
                auto key_at = [&](int pos) { return TMA_KEY ? row[pos * R] : ldg_stream(row + pos * R); };
                if constexpr (PAIR_LD) {
                    uint32_t v0[32], v1[32];
                    tmem_ld32(t_buf + (uint32_t)(rr * 64), v0);
                    tmem_ld32(t_buf + (uint32_t)(rr * 64 + 32), v1);
                    tmem_wait_ld();
                    static_for<0, 8>([&](auto e) {
                        C::mac_one(F[decltype(e)::value], key_at(brev(decltype(e)::value, FFT::LOGR)), cplx_from_words(v0, decltype(e)::value));
                    });
                    static_for<0, 8>([&](auto e) {
                        C::mac_one(F[8 + decltype(e)::value], key_at(brev(8 + decltype(e)::value, FFT::LOGR)), cplx_from_words(v1, decltype(e)::value));
                    });
                } else {
                    static_for<0, 2>([&](auto half) {
                        cplx g[8];
                        static_for<0, 8>([&](auto e) {
                            constexpr int pos = brev(decltype(half)::value * 8 + decltype(e)::value, FFT::LOGR);
                            g[decltype(e)::value] = key_at(pos);
                        });
                        uint32_t v[32];
                        tmem_ld32(t_buf + (uint32_t)(rr * 64 + decltype(half)::value * 32), v);
                        tmem_wait_ld();
                        static_for<0, 8>([&](auto e) {
                            C::mac_one(F[decltype(half)::value * 8 + decltype(e)::value], g[decltype(e)::value], cplx_from_words(v, decltype(e)::value));
                        });
                    });
                }
            }
            if constexpr (TMA_KEY) {
                // release the key block: the LAST of the CTA's warps to finish this level's multiply-accumulate starts the next fetch (nobody
                // blocks on a producer); the fetch then overlaps the next level's decomposition + transform
                __syncwarp();
                if (lane == 0) {
                    __threadfence_block();
                    if (atomicAdd(done_cnt, 1u) == (unsigned)(THREADS / 32 - 1)) {
                        atomicExch(done_cnt, 0u);
                        __threadfence_block();
                        if (step + 1 < total_steps) fetch_level(step + 1);
                    }
                }
            }
            // no second barrier: the next write of THIS buffer is two levels ahead, after the barrier of the next level, which the
            // other teams reach only after these reads
        }
        C::inv_pass1(l16, F, region, twi);
        __syncwarp();
        C::inv_pass2_accumulate(l16, region, acc_r);
        __syncwarp();
    }

    tmem_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(X16_COLS));
    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = l16; c < N; c += R) o[(size_t)r * N + c] = acc_r[c];
    } else {
        if (r < K) {
            for (int t = l16; t < N; t += R) o[(size_t)r * N + t] = (t == 0) ? acc_r[0] : (T)(0 - acc_r[N - t]);
        } else if (l16 == 0) {
            o[(size_t)K * N] = acc_r[0];
        }
    }
#endif
}

// Second organisation of the same kernel ("all transforms first"): the decomposition states of a thread's 32 coefficients are built
// ONCE per iteration, the l forward transforms run back to back with nothing else live (no spill), their outputs go to single-buffered
// tensor-memory slots -- and, when the (k+1) l slots do not fit the 512 columns, the last level goes to the team's own (then idle)
// transpose region in shared memory --, ONE barrier, then the whole multiply-accumulate (level-major, row ascending: the reference's
// order) with F in registers, a second barrier (slots and regions are reused by the next iteration / the inverse transform), inverse
// transform, accumulate.  Two barriers per iteration instead of l.
// NST > 0: the bootstrap key is streamed through a ring of NST shared-memory stages by the TMA engine (cp.async.bulk + mbarrier), one
// (level, row) block = (k+1) M complex values = 12 KB per stage: the fetches run ahead of the multiply-accumulate (also during the
// transforms), the MAC reads the key with LDS.  The LAST warp of the CTA to finish a block refills its stage -- no producer warp.
template <typename T, int K, int L, int BL, int NST>
__global__ void __launch_bounds__(4 * (K + 1) * 32, 1)
blind_rotate_x16b_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                         const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                         size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                         const cplx *__restrict__ tw_inv_) {
#if defined(__CUDA_ARCH__)
    constexpr int R = 16, K1 = K + 1, THREADS = 4 * K1 * 32;
    constexpr int TL = (X16_COLS / 64 / K1) < L ? (X16_COLS / 64 / K1) : L;     // levels whose D live in tensor memory
    static_assert(L - TL <= 1, "at most one level's D goes through the team's shared-memory region");
    typedef BRCore<T, R, K, L, BL> C;
    typedef TeamFFT<R> FFT;
    typedef typename StateWord<T, L * BL>::type S;
    constexpr int N = C::N, M = C::M, REGION = C::REGION;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, r = warp >> 2;
    const int h = lane >> 4, l16 = lane & 15;
    const int slot = 2 * q + h;
    int ct = blockIdx.x * X16_CTS + slot;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    constexpr size_t REG_BYTES = (size_t)REGION * sizeof(cplx), ACC_BYTES = (size_t)N * sizeof(T);
    cplx *regions_ct = reinterpret_cast<cplx *>(smem_raw + (size_t)(slot * K1) * REG_BYTES);     // the K1 regions of this ciphertext
    cplx *region = regions_ct + (size_t)r * REGION;
    T *acc_r = reinterpret_cast<T *>(smem_raw + X16_CTS * K1 * REG_BYTES + (size_t)(slot * K1 + r) * ACC_BYTES);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + X16_CTS * K1 * (REG_BYTES + ACC_BYTES) + slot * ahat_bytes);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + X16_CTS * K1 * (REG_BYTES + ACC_BYTES) + X16_CTS * ahat_bytes);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);
    // key ring: [NST blocks | NST mbarriers | NST counters]
    constexpr uint32_t BLOCK_BYTES = (uint32_t)(K1 * M * sizeof(cplx));
    constexpr int NB = L * K1;                                   // blocks per iteration, in consumption order (level l..1, row ascending)
    unsigned char *ring_raw = reinterpret_cast<unsigned char *>(tw_s + 2 * TW_HALF);
    const cplx *ring = reinterpret_cast<const cplx *>(ring_raw);
    const uint32_t ring_sa = smem_addr_u32(ring_raw), bars_sa = ring_sa + (NST > 0 ? NST : 1) * BLOCK_BYTES;
    unsigned int *done_cnt = reinterpret_cast<unsigned int *>(ring_raw + (size_t)(NST > 0 ? NST : 1) * BLOCK_BYTES + 8 * (NST > 0 ? NST : 1));
    const uint32_t total_blocks = (uint32_t)n * NB;
    auto fetch_block = [&](uint32_t g) {           // one thread: global block g = (iteration g / NB, block g % NB) into stage g % NST
        const uint32_t b = g % NB, st = g % (NST > 0 ? NST : 1);
        const uint32_t t_ = b / K1, rr_ = b % K1;
        const unsigned char *src = reinterpret_cast<const unsigned char *>(bsk + (size_t)(g / NB) * C::GGSW_CPLX + ((size_t)(L - 1 - t_) * K1 + rr_) * K1 * M);
        const uint32_t bar = bars_sa + 8 * st;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(BLOCK_BYTES) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ring_sa + st * BLOCK_BYTES),
                     "l"(src), "r"(BLOCK_BYTES), "r"(bar)
                     : "memory");
    };
    if constexpr (NST > 0) {
        if (threadIdx.x == 0) {
            for (int s_ = 0; s_ < NST; s_++) {
                mbarrier_init(bars_sa + 8 * s_, 1);
                done_cnt[s_] = 0;
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "n"(X16_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = r * R + l16; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    tmem_fence_before();
    __syncthreads();
    tmem_fence_after();
    const uint32_t t_quad = tmem_base_smem + ((uint32_t)(q * 32) << 16);
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = l16; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncwarp();
    const int bar_q = 1 + q;
    if constexpr (NST > 0) {
        if (threadIdx.x == 0)                         // after the __syncthreads above: barriers initialised and visible
            for (uint32_t g = 0; g < (uint32_t)NST && g < total_blocks; g++) fetch_block(g);
    }

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
        {
            // decomposition states of this thread's 2 x 16 coefficients of acc * X^a - acc, once per iteration
            S s0[R], s1[R];
            static_for<0, R>([&](auto p) {
                constexpr int m = brev(decltype(p)::value, FFT::LOGR);
                const int c = l16 + R * m;
                s0[decltype(p)::value] = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c, a) - acc_r[c], BL, L);
                s1[decltype(p)::value] = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M], BL, L);
            });
#pragma unroll
            for (int t = 0; t < L; t++) {             // t = 0 <-> level L, the first digit the recurrence emits
                cplx x[R];
                static_for<0, R>([&](auto p) {
                    x[decltype(p)::value] = cplx{int_to_double((int32_t)decomp_next_digit<S>(s0[decltype(p)::value], BL)),
                                                 int_to_double((int32_t)decomp_next_digit<S>(s1[decltype(p)::value], BL))};
                });
                FFT::fwd_pass1(l16, x, region, twf);
                __syncwarp();
                FFT::fwd_pass2_load(l16, x, region);
                __syncwarp();
                FFT::fwd_pass2_compute(x);            // x[p] = D_r[t] at Fourier position p * R + l16
                if (t < TL) {
                    const uint32_t t_mine = t_quad + (uint32_t)((t * K1 + r) * 64);
                    static_for<0, 2>([&](auto half) {
                        uint32_t v[32];
                        static_for<0, 8>([&](auto e) {
                            constexpr int pos = brev(decltype(half)::value * 8 + decltype(e)::value, FFT::LOGR);
                            v[4 * decltype(e)::value + 0] = (uint32_t)__double2loint(x[pos].re);
                            v[4 * decltype(e)::value + 1] = (uint32_t)__double2hiint(x[pos].re);
                            v[4 * decltype(e)::value + 2] = (uint32_t)__double2loint(x[pos].im);
                            v[4 * decltype(e)::value + 3] = (uint32_t)__double2hiint(x[pos].im);
                        });
                        tmem_st32(t_mine + decltype(half)::value * 32, v);
                    });
                } else {
                    // the last level's D through the team's own region, register order pp (position brev(pp)): [pp][l16], unpadded
                    static_for<0, R>([&](auto pp) { region[decltype(pp)::value * R + l16] = x[brev(decltype(pp)::value, FFT::LOGR)]; });
                }
            }
        }
        tmem_wait_st();
        tmem_fence_before();
        named_bar_sync(bar_q, K1 * 32);
        tmem_fence_after();
        // F_r = sum over levels (l down to 1 = t ascending) and rows (ascending) of G[level][row][r] (.) D[row][t]: the reference's order
        cplx F[R];
        static_for<0, R>([&](auto pp) { F[decltype(pp)::value] = cplx{0.0, 0.0}; });
        // software-pipelined over the 2 l (k+1) half-rows: the key values of half-row c + 1 are requested from L2 before half-row c is
        // consumed (the three warps of a scheduler are all in this phase together: nobody else hides the latency)
        constexpr int NCH = 2 * L * K1;
        const cplx *gcol = ggsw_i + (size_t)r * M + l16;
        auto key_load = [&](auto c, cplx (&g)[8]) {
            constexpr int t = decltype(c)::value / (2 * K1), rr = (decltype(c)::value / 2) % K1, half = decltype(c)::value % 2;
            const cplx *row = gcol + ((size_t)(L - 1 - t) * K1 + rr) * K1 * M;
            static_for<0, 8>([&](auto e) { g[decltype(e)::value] = ldg_stream(row + brev(half * 8 + decltype(e)::value, FFT::LOGR) * R); });
        };
        cplx ga[8], gb[8];
        if constexpr (NST == 0) key_load(std::integral_constant<int, 0>{}, ga);
        static_for<0, NCH>([&](auto c) {
            constexpr int t = decltype(c)::value / (2 * K1), rr = (decltype(c)::value / 2) % K1, half = decltype(c)::value % 2;
            cplx(&g)[8] = (decltype(c)::value % 2 == 0) ? ga : gb;
            cplx(&gn)[8] = (decltype(c)::value % 2 == 0) ? gb : ga;
            if constexpr (NST == 0) {
                if constexpr (decltype(c)::value + 1 < NCH) key_load(std::integral_constant<int, decltype(c)::value + 1>{}, gn);
            } else {
                // block gblk = i * NB + c / 2 sits in stage gblk % NST; its phase parity is (gblk / NST) & 1
                const uint32_t gblk = (uint32_t)i * NB + (uint32_t)(decltype(c)::value / 2), st = gblk % NST;
                if constexpr (half == 0) mbarrier_wait(bars_sa + 8 * st, (gblk / NST) & 1);
                const cplx *kr = ring + (size_t)st * (K1 * M) + (size_t)r * M + l16;
                static_for<0, 8>([&](auto e) { g[decltype(e)::value] = kr[brev(half * 8 + decltype(e)::value, FFT::LOGR) * R]; });
            }
            if constexpr (t < TL) {
                uint32_t v[32];
                tmem_ld32(t_quad + (uint32_t)((t * K1 + rr) * 64 + half * 32), v);
                tmem_wait_ld();
                static_for<0, 8>([&](auto e) { C::mac_one(F[half * 8 + decltype(e)::value], g[decltype(e)::value], cplx_from_words(v, decltype(e)::value)); });
            } else {
                const cplx *d = regions_ct + (size_t)rr * REGION + l16;
                static_for<0, 8>([&](auto e) {
                    constexpr int pp = half * 8 + decltype(e)::value;
                    C::mac_one(F[pp], g[decltype(e)::value], d[pp * R]);
                });
            }
            if constexpr (NST > 0 && half == 1) {
                // this warp is done with the block: the last of the CTA's warps refills the stage with the block NST ahead
                const uint32_t gblk = (uint32_t)i * NB + (uint32_t)(decltype(c)::value / 2), st = gblk % NST;
                __syncwarp();
                if (lane == 0) {
                    __threadfence_block();
                    if (atomicAdd(&done_cnt[st], 1u) == (unsigned)(THREADS / 32 - 1)) {
                        atomicExch(&done_cnt[st], 0u);
                        __threadfence_block();
                        if (gblk + NST < total_blocks) fetch_block(gblk + NST);
                    }
                }
            }
        });
        FFT::inv_pass1_compute(l16, F, twi);     // butterflies and middle twiddle BEFORE the barrier: only the store needs the region free
        tmem_fence_before();
        named_bar_sync(bar_q, K1 * 32);          // every team is done with the D slots and regions of this iteration
        tmem_fence_after();
        FFT::pass1_store(l16, F, region);
        __syncwarp();
        C::inv_pass2_accumulate(l16, region, acc_r);
        __syncwarp();
    }

    tmem_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(X16_COLS));
    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = l16; c < N; c += R) o[(size_t)r * N + c] = acc_r[c];
    } else {
        if (r < K) {
            for (int t = l16; t < N; t += R) o[(size_t)r * N + t] = (t == 0) ? acc_r[0] : (T)(0 - acc_r[N - t]);
        } else if (l16 == 0) {
            o[(size_t)K * N] = acc_r[0];
        }
    }
#endif
}

// boolean DEFAULT_PARAMETERS (boolean/parameters/mod.rs:123-134): u32, k 2, N 512, level 3, base 2^6
#define X16_INSTANCES(X) X(uint32_t, 2, 3, 6)

int br_x16_supported(const Params &p) {
#define X(T, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 512 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return 16;
    X16_INSTANCES(X)
#undef X
    return 0;
}

static int g_x16_variant = 12;     // default: "all transforms first" + TMA key ring (the fastest measured); 0..3 select the first organisation
void br_x16_set_variant(int v) { g_x16_variant = v < 0 ? 12 : (v & 15); }

template <typename T, int K, int L, int BL, int VAR> static int launch_x16v(const BRLaunch &a, cudaStream_t st) {
    typedef BRCore<T, 16, K, L, BL> C;
    constexpr int K1 = K + 1;
    auto kern = blind_rotate_x16_kernel<T, K, L, BL, VAR>;
    const size_t ahat = (((size_t)(a.n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    size_t smem = (size_t)X16_CTS * K1 * ((size_t)C::REGION * sizeof(cplx) + (size_t)C::N * sizeof(T)) + X16_CTS * ahat + sizeof(TwTables<16>);
    if (VAR & 1) smem += (size_t)K1 * K1 * C::M * sizeof(cplx) + 16;      // one level of the key + mbarrier + counter
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    const int grid = (a.batch + X16_CTS - 1) / X16_CTS;
    kern<<<grid, 4 * K1 * 32, smem, st>>>((const T *)a.lwe_in, a.lwe_stride, (const T *)a.luts, a.lut_idx, (const cplx *)a.bsk, (T *)a.out,
                                          a.out_stride, a.n, a.batch, a.mode, (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv);
    return (int)cudaGetLastError();
}
template <typename T, int K, int L, int BL, int NST> static int launch_x16b(const BRLaunch &a, cudaStream_t st) {
    typedef BRCore<T, 16, K, L, BL> C;
    constexpr int K1 = K + 1;
    auto kern = blind_rotate_x16b_kernel<T, K, L, BL, NST>;
    const size_t ahat = (((size_t)(a.n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    const size_t smem = (size_t)X16_CTS * K1 * ((size_t)C::REGION * sizeof(cplx) + (size_t)C::N * sizeof(T)) + X16_CTS * ahat + sizeof(TwTables<16>) +
                        (NST > 0 ? (size_t)NST * (K1 * C::M * sizeof(cplx) + 16) : 0);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    const int grid = (a.batch + X16_CTS - 1) / X16_CTS;
    kern<<<grid, 4 * K1 * 32, smem, st>>>((const T *)a.lwe_in, a.lwe_stride, (const T *)a.luts, a.lut_idx, (const cplx *)a.bsk, (T *)a.out,
                                          a.out_stride, a.n, a.batch, a.mode, (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv);
    return (int)cudaGetLastError();
}
template <typename T, int K, int L, int BL> static int launch_x16(const BRLaunch &a, cudaStream_t st) {
    if ((g_x16_variant & 12) == 12) return launch_x16b<T, K, L, BL, 4>(a, st);
    if (g_x16_variant & 4) return launch_x16b<T, K, L, BL, 0>(a, st);
    switch (g_x16_variant) {
        case 1: return launch_x16v<T, K, L, BL, 1>(a, st);
        case 2: return launch_x16v<T, K, L, BL, 2>(a, st);
        case 3: return launch_x16v<T, K, L, BL, 3>(a, st);
        default: return launch_x16v<T, K, L, BL, 0>(a, st);
    }
}

int br_x16_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (a.batch <= 0) return 0;
#define X(T, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 512 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return launch_x16<T, K, L, BL>(a, st);
    X16_INSTANCES(X)
#undef X
    return -1;
}

}  // namespace b200tfhe

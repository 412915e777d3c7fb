// blind_rotate_lat.cu -- single-ciphertext ("latency") blind rotation for N = 2048, k = 1, l = 1 (PARAM_MESSAGE_2_CARRY_2).
//
// The batch kernels give one polynomial to one warp (32 points per thread) and hide every latency behind the other
// ciphertexts of the SM.  With ONE ciphertext per SM that leaves two warps on two of the four schedulers, and ncu shows
// the iteration bound by instruction issue: 5000 SASS instructions per warp-iteration, 1.1 "no instruction" stall cycles
// per issue because the unrolled 80 KB loop body misses the instruction cache with nobody to share the fetch
// (profiles/r01_ncu_latency.json).  This kernel is built for that regime instead:
//   * 128 threads per polynomial, 8 points each (fft8.cuh: 1024 = 8 x 16 x 8), 8 warps = two per scheduler, one of each
//     polynomial: the shared-memory exchanges of one overlap the butterflies of the other (round 1: 64 threads x 16 points, one
//     warp per scheduler, shared-memory pipe 66 % + FP64 pipe 38 % busy strictly one after the other);
//   * the forward and the inverse transform share ONE copy of the FFT code: the inverse is the forward of the conjugate,
//     the loop over the two directions is not unrolled, direction-dependent work sits in uniform branches around it;
//   * this iteration's Fourier GGSW (64 KB) is staged into shared memory by the TMA engine (cp.async.bulk + mbarrier)
//     while the previous inverse transform and the next decomposition run: no key registers, no exposed L2 latency;
//   * the closed-form rotated digit (tmem_util.cuh) replaces the generic decomposition recurrence.
// Same mathematics and same bootstrap-key layout as blind_rotate.cu (reference: fft64/crypto/bootstrap.rs:243-412,
// fft64/crypto/ggsw.rs:477-598); the stage sequence is executed on the CPU by tests/native/emul.cpp.
#include "fft8.cuh"
#include "tmem_util.cuh"
#include "kernels.h"
#include "twiddles.hpp"
#include <map>
#include <mutex>

namespace b200tfhe {

namespace {

constexpr int LAT_TEAM = Fft8::T, LAT_THREADS = 2 * LAT_TEAM, LAT_N = 2048, LAT_M = 1024, LAT_K1 = 2, LAT_BL = 23;
constexpr size_t LAT_KEY_BYTES = (size_t)LAT_K1 * LAT_K1 * LAT_M * sizeof(cplx);        // 65536: one Fourier GGSW
constexpr size_t LAT_ACC_BYTES = (size_t)LAT_K1 * LAT_N * sizeof(uint64_t);             // 32768
constexpr size_t LAT_BUF1_BYTES = (size_t)LAT_K1 * Fft8::BUF1 * sizeof(cplx);           // 32768
constexpr size_t LAT_BUF2_BYTES = (size_t)LAT_K1 * Fft8::BUF2 * sizeof(cplx);           // 36864
constexpr size_t LAT_OFF_ACC = LAT_KEY_BYTES;
constexpr size_t LAT_OFF_BUF1 = LAT_OFF_ACC + LAT_ACC_BYTES;
constexpr size_t LAT_OFF_BUF2 = LAT_OFF_BUF1 + LAT_BUF1_BYTES;
constexpr size_t LAT_OFF_TAB = LAT_OFF_BUF2 + LAT_BUF2_BYTES;
constexpr size_t LAT_OFF_MBAR = LAT_OFF_TAB + sizeof(Fft8::Tables);
constexpr size_t LAT_OFF_AHAT = LAT_OFF_MBAR + 16;
static_assert(sizeof(Fft8::Tables) % 16 == 0, "table block keeps 16-byte alignment");

size_t lat_smem_bytes(int n) { return LAT_OFF_AHAT + ((((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15); }

#if defined(__CUDA_ARCH__)
DEV void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
DEV void mbarrier_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
DEV cplx shfl_xor_cplx(cplx v, int mask) {
    return cplx{__shfl_xor_sync(0xffffffffu, v.re, mask), __shfl_xor_sync(0xffffffffu, v.im, mask)};
}
#endif

// BL: base log of the one-level decomposition.  PLAIN = false: the classic CMUX iteration acc += GGSW_i (.) (acc X^{a_i} - acc), n_iter = n.
// PLAIN = true: acc <- GGSW'_g (.) acc, the iteration of the multi-bit blind rotation once the monomial-weighted sum GGSW'_g of a
// group's key ciphertexts has been formed (blind_rotate_mb.cu: mb_key_sum_kernel); n_iter = n / grouping, the ciphertext's own list
// of summed keys starts at bsk + ct * key_ct_stride.  Without the rotated reads the accumulator lives in registers only.
template <int BL, bool PLAIN>
__global__ void __launch_bounds__(LAT_THREADS, 1)
blind_rotate_lat_kernel(const uint64_t *__restrict__ lwe_in, size_t lwe_stride, const uint64_t *__restrict__ luts,
                        const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, uint64_t *__restrict__ out,
                        size_t out_stride, int n, int mode, const Fft8::Tables *__restrict__ tables_g, int n_iter,
                        size_t key_ct_stride) {
#if defined(__CUDA_ARCH__)
    typedef uint64_t T;
    constexpr int N = LAT_N, M = LAT_M, K1 = LAT_K1, TM = LAT_TEAM;
    extern __shared__ __align__(128) unsigned char smem[];
    cplx *key = reinterpret_cast<cplx *>(smem);                               // [r][j][M], refilled every iteration by TMA
    T *acc = reinterpret_cast<T *>(smem + LAT_OFF_ACC);                       // [K1][N]
    cplx *buf1 = reinterpret_cast<cplx *>(smem + LAT_OFF_BUF1);               // [K1][BUF1]
    cplx *buf2 = reinterpret_cast<cplx *>(smem + LAT_OFF_BUF2);               // [K1][BUF2]
    Fft8::Tables *tab = reinterpret_cast<Fft8::Tables *>(smem + LAT_OFF_TAB);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem + LAT_OFF_AHAT);
    const uint32_t bar = smem_addr_u32(smem + LAT_OFF_MBAR);
    const uint32_t key_s = smem_addr_u32(key);
    __shared__ uint32_t tmem_base_smem;

    const int tid = threadIdx.x, r = tid >> 7, t = tid & (TM - 1);
    if (tid < 32) {   // 64 tensor-memory columns: the two teams hand each other their Fourier polynomial D_r through them
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr_u32(&tmem_base_smem)), "n"(64));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tmem_fence_before();
    const int ct = blockIdx.x;
    bsk += (size_t)ct * key_ct_stride;
    T *acc_r = acc + r * N;
    cplx *buf1_r = buf1 + r * Fft8::BUF1, *buf2_r = buf2 + r * Fft8::BUF2;

    if (tid == 0) {
        mbarrier_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    {   // twiddle tables: global -> shared, once
        const int4 *src = reinterpret_cast<const int4 *>(tables_g);
        int4 *dst = reinterpret_cast<int4 *>(tab);
        for (int e = tid; e < (int)(sizeof(Fft8::Tables) / 16); e += LAT_THREADS) dst[e] = src[e];
    }
    // modulus switch of the whole input (fft_impl/common.rs:26-43)
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = tid; i <= n; i += LAT_THREADS) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], 11) & (2 * N - 1));
    __syncthreads();
    tmem_fence_after();
    // thread (r, t) and thread (1 - r, t) sit in warps w and w + 4: same lane quadrant, same lane, so what one stores in
    // columns [32 r, 32 r + 32) the other can load (32x32b: a thread reaches only its own lane)
    const uint32_t t_lane = tmem_base_smem + ((uint32_t)(((tid >> 5) & 3) * 32) << 16);
    const uint32_t t_own = t_lane + (uint32_t)(32 * r), t_oth = t_lane + (uint32_t)(32 * (1 - r));
    auto fetch_key = [&](int i) {   // thread 0: this iteration's GGSW, four 16 KB rows
        mbarrier_expect_tx(bar, (uint32_t)LAT_KEY_BYTES);
        const unsigned char *src = reinterpret_cast<const unsigned char *>(bsk) + (size_t)i * LAT_KEY_BYTES;
#pragma unroll
        for (int q = 0; q < 4; q++) bulk_load(key_s + q * 16384u, src + q * 16384u, 16384u, bar);
    };
    if (tid == 0 && n_iter > 0) fetch_key(0);
    {   // acc <- LUT * X^{-b_hat}  (polynomial_algorithms.rs:315-354)
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = t; c < N; c += TM) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncthreads();

    const unsigned char *accb = reinterpret_cast<const unsigned char *>(acc_r);
    const cplx *key_own = key + (size_t)(r * K1 + r) * M + t;            // row r (own transform), column j = r
    const cplx *key_oth = key + (size_t)((1 - r) * K1 + r) * M + t;      // row 1-r, column j = r

    // the 16 accumulator coefficients this thread rounds into (t + 128 q, and + M) stay in registers from one
    // accumulation to the next decomposition; shared memory keeps the copy the rotated reads need
    // this thread's twiddle factors (fft8.cuh), in registers for the whole blind rotation
    const cplx twist_t = tab->twist[t];
    const Fft8::TwB twb_fwd = tab->tb[0][t], twb_inv = tab->tb[1][t];
    T own_re[8], own_im[8];
    static_for<0, 8>([&](auto q) {
        own_re[decltype(q)::value] = acc_r[t + TM * decltype(q)::value];
        own_im[decltype(q)::value] = acc_r[t + TM * decltype(q)::value + M];
    });

    for (int i = 0; i < n_iter; i++) {
        const int a = PLAIN ? 0 : ahat[i];
        cplx x[8];
#pragma unroll
        for (int dir = 0; dir < 2; dir++) {
            if (dir == 0) {
                // rotate-subtract + one-level signed decomposition, as doubles; then the register part of the twist
                if constexpr (PLAIN) {
                    typedef BRCore<T, 32, 1, 1, BL> DC;
                    static_for<0, 8>([&](auto p) {
                        constexpr int v = brev(decltype(p)::value, 3);
                        x[decltype(p)::value].re = int_to_double(DC::single_level_digit(own_re[v]));
                        x[decltype(p)::value].im = int_to_double(DC::single_level_digit(own_im[v]));
                    });
                } else {
                    const uint32_t s8 = 8u * (uint32_t)(t - a);
                    static_for<0, 8>([&](auto p) {
                        constexpr int v = brev(decltype(p)::value, 3);
                        x[decltype(p)::value].re = rot_digit_f64<BL>(accb, s8 + 8u * (uint32_t)TM * v, own_re[v]);
                        x[decltype(p)::value].im = rot_digit_f64<BL>(accb, s8 + 8u * (uint32_t)TM * v + 8u * M, own_im[v]);
                    });
                }
                Fft8::pre_twist(x);
            } else {
                // x holds D_r[t + 128 f2]; publish it, then F_j = G[r][j] D_r + G[1-r][j] D_{1-r} for j = r, conjugated, in
                // the register order pass A wants (element brev3(p))
                tmem_store_cplx8(t_own, x[0], x[1], x[2], x[3], x[4], x[5], x[6], x[7]);
                tmem_wait_st();
                tmem_fence_before();
                __syncthreads();
                tmem_fence_after();
                uint32_t dw[32];
                tmem_ld32(t_oth, dw);
                mbarrier_wait(bar, (uint32_t)(i & 1));
                cplx y[8];
                static_for<0, 8>([&](auto p) {       // own row first: needs no tensor-memory data
                    constexpr int v = brev(decltype(p)::value, 3);
                    y[decltype(p)::value] = cmul(key_own[TM * v], x[v]);
                });
                tmem_wait_ld();
                static_for<0, 8>([&](auto p) {
                    constexpr int v = brev(decltype(p)::value, 3);
                    cplx f = y[decltype(p)::value];
                    BRCore<T, 32, 1, 1, BL>::mac_one(f, key_oth[TM * v], cplx_from_words(dw, v));
                    x[decltype(p)::value] = cplx{f.re, -f.im};
                });
                tmem_fence_before();
                __syncthreads();   // both teams are done with D (tensor memory) and with the key
                tmem_fence_after();
                if (tid == 0 && i + 1 < n_iter) fetch_key(i + 1);
            }
            Fft8::pass_a(t, x, tab->ta, twist_t, dir == 0, buf1_r);
            named_bar_sync(1 + r, TM);
            {
                cplx z[8], send[4], recv[4];
                Fft8::pass_b_first(t, buf1_r, z, send);
                static_for<0, 4>([&](auto e) { recv[decltype(e)::value] = shfl_xor_cplx(send[decltype(e)::value], 8); });
                Fft8::pass_b_second(t, z, recv, dir ? twb_inv : twb_fwd, buf2_r);
            }
            named_bar_sync(1 + r, TM);
            Fft8::pass_c(t, x, buf2_r);
            if (dir == 1) {
                Fft8::post_twist(x);
                if constexpr (PLAIN) {
                    static_for<0, 8>([&](auto q) {
                        own_re[decltype(q)::value] = torus_from_double<T>(x[decltype(q)::value].re);
                        own_im[decltype(q)::value] = torus_from_double<T>(-x[decltype(q)::value].im);
                    });
                } else {
                    static_for<0, 8>([&](auto q) {
                        const int c = t + TM * decltype(q)::value;
                        own_re[decltype(q)::value] += torus_from_double<T>(x[decltype(q)::value].re);
                        own_im[decltype(q)::value] += torus_from_double<T>(-x[decltype(q)::value].im);
                        acc_r[c] = own_re[decltype(q)::value];
                        acc_r[c + M] = own_im[decltype(q)::value];
                    });
                    named_bar_sync(1 + r, TM);
                }
            }
        }
    }
    if constexpr (PLAIN) {
        static_for<0, 8>([&](auto q) {
            acc_r[t + TM * decltype(q)::value] = own_re[decltype(q)::value];
            acc_r[t + TM * decltype(q)::value + M] = own_im[decltype(q)::value];
        });
        named_bar_sync(1 + r, TM);
    }

    tmem_fence_before();
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(64));
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = t; c < N; c += TM) o[(size_t)r * N + c] = acc_r[c];
    } else if (r == 0) {   // sample extract, coefficient 0 (algorithms/glwe_sample_extraction.rs:91-147), k = 1
        for (int c = t; c < N; c += TM) o[c] = (c == 0) ? acc_r[0] : (T)(0 - acc_r[N - c]);
    } else if (t == 0) {
        o[N] = acc_r[0];
    }
#endif
}

// ---- one ciphertext on a CLUSTER of two CTAs (two SMs): CTA r owns polynomial row r -------------------------------------------------
// Each CTA runs one 128-thread team (one warp per scheduler) with its own accumulator row, exchange buffers and the key COLUMN
// j = r (two 16 KB blocks per iteration, by TMA, two iterations ahead).  The only data that crosses between the SMs is the transformed
// polynomial D_r: every thread sends its eight values into the PEER's shared memory with st.async (distributed shared memory), each store
// signalling the peer's mbarrier; the landing buffer is double-buffered by iteration parity and there is no cluster barrier in the loop.
// Half the shared-memory traffic and half the FP64 work per SM; the price is the 16 KB that cross at ~20 B/clk.  Used for batches of at
// most `lat_cluster_max` ciphertexts (two SMs each).  History of the exchange: remote stores + cluster barrier 2.06 ms,
// one staged 16 KB bulk copy 1.80 ms, st.async 1.71 ms of blind rotation.  -DLAT2_EXP_NO_SEND / -DLAT2_EXP_NO_DIN_WAIT build a kernel without
// the exchange (WRONG results; timing experiments only: 1.385 ms, i.e. the exchange costs ~855 cycles of the 4.5 k of an iteration).
constexpr size_t L2_KEY_BYTES = LAT_KEY_BYTES / 2;                                   // column j = r: rows 0 and 1
constexpr size_t L2_OFF_ACC = 2 * L2_KEY_BYTES;                                       // two key slots: the fetch runs two iterations ahead
constexpr size_t L2_OFF_BUF1 = L2_OFF_ACC + (size_t)LAT_N * sizeof(uint64_t);
constexpr size_t L2_OFF_BUF2 = L2_OFF_BUF1 + (size_t)Fft8::BUF1 * sizeof(cplx);
constexpr size_t L2_OFF_DIN = L2_OFF_BUF2 + (size_t)Fft8::BUF2 * sizeof(cplx);       // [2][M] landing buffers for the peer's D
constexpr size_t L2_OFF_TAB = L2_OFF_DIN + 2 * (size_t)LAT_M * sizeof(cplx);
constexpr size_t L2_OFF_MBAR = L2_OFF_TAB + sizeof(Fft8::Tables);
constexpr size_t L2_OFF_AHAT = L2_OFF_MBAR + 32;   // three mbarriers: key slot 0, key slot 1, landing buffer
size_t lat2_smem_bytes(int n) { return L2_OFF_AHAT + ((((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15); }

template <int BL, bool PLAIN>
__global__ void __launch_bounds__(LAT_TEAM, 1)
blind_rotate_lat2_kernel(const uint64_t *__restrict__ lwe_in, size_t lwe_stride, const uint64_t *__restrict__ luts,
                         const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, uint64_t *__restrict__ out,
                         size_t out_stride, int n, int mode, const Fft8::Tables *__restrict__ tables_g, int n_iter,
                         size_t key_ct_stride) {
#if defined(__CUDA_ARCH__)
    typedef uint64_t T;
    constexpr int N = LAT_N, M = LAT_M, K1 = LAT_K1, TM = LAT_TEAM;
    extern __shared__ __align__(128) unsigned char smem[];
    cplx *key = reinterpret_cast<cplx *>(smem);                               // [row q][M]: G[q][j = r], refilled every iteration by TMA
    T *acc_r = reinterpret_cast<T *>(smem + L2_OFF_ACC);                      // this CTA's accumulator row
    cplx *buf1 = reinterpret_cast<cplx *>(smem + L2_OFF_BUF1);
    cplx *buf2 = reinterpret_cast<cplx *>(smem + L2_OFF_BUF2);
    cplx *din = reinterpret_cast<cplx *>(smem + L2_OFF_DIN);
    Fft8::Tables *tab = reinterpret_cast<Fft8::Tables *>(smem + L2_OFF_TAB);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem + L2_OFF_AHAT);
    const uint32_t bar = smem_addr_u32(smem + L2_OFF_MBAR);
    const uint32_t key_s = smem_addr_u32(key);

    const int t = threadIdx.x;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int r = (int)rank, ct = blockIdx.x >> 1;
    bsk += (size_t)ct * key_ct_stride;
    // the peer's landing buffer and its "D has landed" mbarrier, as shared::cluster addresses
    const uint32_t bar_din = bar + 16;
    uint32_t din_peer, bar_din_peer;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(din_peer) : "r"(smem_addr_u32(din)), "r"(rank ^ 1u));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(bar_din_peer) : "r"(bar_din), "r"(rank ^ 1u));

    if (t == 0) {
        mbarrier_init(bar, 1);
        mbarrier_init(bar + 8, 1);
        mbarrier_init(bar_din, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    {
        const int4 *src = reinterpret_cast<const int4 *>(tables_g);
        int4 *dst = reinterpret_cast<int4 *>(tab);
        for (int e = t; e < (int)(sizeof(Fft8::Tables) / 16); e += TM) dst[e] = src[e];
    }
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = t; i <= n; i += TM) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], 11) & (2 * N - 1));
    __syncthreads();
    // thread 0: column j = r of iteration i's GGSW (rows 0 and 1, 16 KB each) into key slot i & 1.  The fetch runs TWO iterations ahead
    // of its use, so that it is covered even when the key comes from HBM (a keyswitch in front of the bootstrap streams 61 MB through L2)
    auto fetch_key = [&](int i) {
        const uint32_t kb = bar + 8u * (uint32_t)(i & 1), ks = key_s + (uint32_t)((i & 1) * L2_KEY_BYTES);
        mbarrier_expect_tx(kb, (uint32_t)L2_KEY_BYTES);
        const unsigned char *src = reinterpret_cast<const unsigned char *>(bsk) + (size_t)i * LAT_KEY_BYTES;
#pragma unroll
        for (int q = 0; q < 2; q++) bulk_load(ks + q * 16384u, src + (size_t)(q * K1 + r) * 16384u, 16384u, kb);
    };
    if (t == 0 && n_iter > 0) fetch_key(0);
    if (t == 0 && n_iter > 1) fetch_key(1);
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = t; c < N; c += TM) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncthreads();
    // both CTAs are up (their shared memory exists) before anyone stores into the peer
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");

    const unsigned char *accb = reinterpret_cast<const unsigned char *>(acc_r);
    const cplx *key_own0 = key + (size_t)r * M + t;            // row r, column r       (slot 0; slot 1 is 2 M further)
    const cplx *key_oth0 = key + (size_t)(1 - r) * M + t;      // row 1 - r, column r
    const cplx twist_t = tab->twist[t];
    const Fft8::TwB twb_fwd = tab->tb[0][t], twb_inv = tab->tb[1][t];
    T own_re[8], own_im[8];
    static_for<0, 8>([&](auto q) {
        own_re[decltype(q)::value] = acc_r[t + TM * decltype(q)::value];
        own_im[decltype(q)::value] = acc_r[t + TM * decltype(q)::value + M];
    });

    for (int i = 0; i < n_iter; i++) {
        const int a = PLAIN ? 0 : ahat[i];
        cplx x[8];
#pragma unroll
        for (int dir = 0; dir < 2; dir++) {
            if (dir == 0) {
                if constexpr (PLAIN) {
                    typedef BRCore<T, 32, 1, 1, BL> DC;
                    static_for<0, 8>([&](auto p) {
                        constexpr int v = brev(decltype(p)::value, 3);
                        x[decltype(p)::value].re = int_to_double(DC::single_level_digit(own_re[v]));
                        x[decltype(p)::value].im = int_to_double(DC::single_level_digit(own_im[v]));
                    });
                } else {
                    const uint32_t s8 = 8u * (uint32_t)(t - a);
                    static_for<0, 8>([&](auto p) {
                        constexpr int v = brev(decltype(p)::value, 3);
                        x[decltype(p)::value].re = rot_digit_f64<BL>(accb, s8 + 8u * (uint32_t)TM * v, own_re[v]);
                        x[decltype(p)::value].im = rot_digit_f64<BL>(accb, s8 + 8u * (uint32_t)TM * v + 8u * M, own_im[v]);
                    });
                }
                Fft8::pre_twist(x);
            } else {
                // x holds D_r[t + 128 f2].  Every thread sends its eight values into the peer's landing buffer with st.async: a remote store
                // through distributed shared memory that signals the PEER's mbarrier with the bytes it delivered.  The sender never waits
                // (no fence, no staging, no barrier); the receiver waits for 16 KB on its own mbarrier.  No cluster barrier in the loop: "the
                // peer's D of iteration i+1 has landed" implies the peer has consumed what this CTA sent in iteration i, so the two landing
                // slots are safe to reuse two iterations later.  The same chain keeps the phases of the landing barrier apart: the peer issues
                // its stores of iteration i+1 only after it has consumed this CTA's D of iteration i, a whole iteration (> 2 us) after its own
                // stores of iteration i left -- they cannot be counted into the wrong phase.
#ifndef LAT2_EXP_NO_SEND
                if (t == 0) mbarrier_expect_tx(bar_din, (uint32_t)(M * sizeof(cplx)));   // arm this CTA's own landing barrier
#endif
#ifndef LAT2_EXP_NO_SEND
                {
                    const uint32_t dst = din_peer + (uint32_t)(((i & 1) * M + t) * sizeof(cplx));
                    static_for<0, 8>([&](auto f2) {
                        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];" ::"r"(
                                         dst + (uint32_t)(TM * decltype(f2)::value * sizeof(cplx))),
                                     "d"(x[decltype(f2)::value].re), "d"(x[decltype(f2)::value].im), "r"(bar_din_peer)
                                     : "memory");
                    });
                }
#endif
                const cplx *key_own = key_own0 + (i & 1) * 2 * M, *key_oth = key_oth0 + (i & 1) * 2 * M;
                mbarrier_wait(bar + 8u * (uint32_t)(i & 1), (uint32_t)((i >> 1) & 1));
                cplx y[8];
                static_for<0, 8>([&](auto p) {       // own row first: needs nothing from the peer
                    constexpr int v = brev(decltype(p)::value, 3);
                    y[decltype(p)::value] = cmul(key_own[TM * v], x[v]);
                });
#ifndef LAT2_EXP_NO_DIN_WAIT
                mbarrier_wait(bar_din, (uint32_t)(i & 1));
#endif
                const cplx *d_oth = din + (i & 1) * M + t;
                static_for<0, 8>([&](auto p) {
                    constexpr int v = brev(decltype(p)::value, 3);
                    cplx f = y[decltype(p)::value];
                    BRCore<T, 32, 1, 1, BL>::mac_one(f, key_oth[TM * v], d_oth[TM * v]);
                    x[decltype(p)::value] = cplx{f.re, -f.im};
                });
            }
            Fft8::pass_a(t, x, tab->ta, twist_t, dir == 0, buf1);
            __syncthreads();
            // every thread of the CTA is past its key reads: refill the key column for the next iteration
            if (dir == 1 && t == 0 && i + 2 < n_iter) fetch_key(i + 2);
            {
                cplx z[8], send[4], recv[4];
                Fft8::pass_b_first(t, buf1, z, send);
                static_for<0, 4>([&](auto e) { recv[decltype(e)::value] = shfl_xor_cplx(send[decltype(e)::value], 8); });
                Fft8::pass_b_second(t, z, recv, dir ? twb_inv : twb_fwd, buf2);
            }
            __syncthreads();
            Fft8::pass_c(t, x, buf2);
            if (dir == 1) {
                Fft8::post_twist(x);
                if constexpr (PLAIN) {
                    static_for<0, 8>([&](auto q) {
                        own_re[decltype(q)::value] = torus_from_double<T>(x[decltype(q)::value].re);
                        own_im[decltype(q)::value] = torus_from_double<T>(-x[decltype(q)::value].im);
                    });
                } else {
                    static_for<0, 8>([&](auto q) {
                        const int c = t + TM * decltype(q)::value;
                        own_re[decltype(q)::value] += torus_from_double<T>(x[decltype(q)::value].re);
                        own_im[decltype(q)::value] += torus_from_double<T>(-x[decltype(q)::value].im);
                        acc_r[c] = own_re[decltype(q)::value];
                        acc_r[c + M] = own_im[decltype(q)::value];
                    });
                    __syncthreads();
                }
            }
        }
    }
    if constexpr (PLAIN) {
        static_for<0, 8>([&](auto q) {
            acc_r[t + TM * decltype(q)::value] = own_re[decltype(q)::value];
            acc_r[t + TM * decltype(q)::value + M] = own_im[decltype(q)::value];
        });
        __syncthreads();
    }
    // no CTA of the cluster exits while the other may still store into it
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");

    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = t; c < N; c += TM) o[(size_t)r * N + c] = acc_r[c];
    } else if (r == 0) {   // sample extract, coefficient 0 (algorithms/glwe_sample_extraction.rs:91-147), k = 1
        for (int c = t; c < N; c += TM) o[c] = (c == 0) ? acc_r[0] : (T)(0 - acc_r[N - c]);
    } else if (t == 0) {
        o[N] = acc_r[0];
    }
#endif
}

// per-device copy of the twiddle tables, uploaded on first use
const Fft8::Tables *lat_tables(cudaStream_t st, cudaError_t *err) {
    static std::mutex mu;
    static std::map<int, Fft8::Tables *> per_device;
    static Fft8::Tables host;
    static bool filled = false;
    int dev = 0;
    if ((*err = cudaGetDevice(&dev)) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> g(mu);
    auto it = per_device.find(dev);
    if (it != per_device.end()) return it->second;
    if (!filled) { fill_fft8_tables(host); filled = true; }
    Fft8::Tables *d = nullptr;
    if ((*err = cudaMalloc(&d, sizeof(Fft8::Tables))) != cudaSuccess) return nullptr;
    if ((*err = cudaMemcpyAsync(d, &host, sizeof(Fft8::Tables), cudaMemcpyHostToDevice, st)) != cudaSuccess ||
        (*err = cudaStreamSynchronize(st)) != cudaSuccess) {
        cudaFree(d);
        return nullptr;
    }
    per_device[dev] = d;
    return d;
}

}  // namespace

template <int BL, bool PLAIN>
static int launch_lat2(const BRLaunch &a, const Fft8::Tables *tab, int n_iter, size_t key_ct_stride, cudaStream_t st) {
    auto kern = blind_rotate_lat2_kernel<BL, PLAIN>;
    const size_t smem = lat2_smem_bytes(a.n);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (unsigned)a.batch);
    cfg.blockDim = dim3(LAT_TEAM);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&cfg, kern, (const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx, (const cplx *)a.bsk,
                                   (uint64_t *)a.out, a.out_stride, a.n, a.mode, tab, n_iter, key_ct_stride);
}

int br_lat_supported(const Params &p) {
    return p.bits == 64 && p.N == LAT_N && p.k == 1 && p.pbs_level == 1 && p.pbs_base_log == LAT_BL;
}
// base logs of PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_{2,3}_KS_PBS (shortint/parameters/multi_bit.rs:115-132,173-190)
#define LAT_PLAIN_INSTANCES(X) X(22) X(21)
int br_lat_plain_supported(const Params &p) {
    if (!(p.bits == 64 && p.N == LAT_N && p.k == 1 && p.pbs_level == 1)) return 0;
#define X(BL) \
    if (p.pbs_base_log == BL) return 1;
    LAT_PLAIN_INSTANCES(X)
#undef X
    return 0;
}

int br_lat_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (!br_lat_supported(p)) return -1;
    if (a.batch <= 0) return 0;
    cudaError_t e = cudaSuccess;
    const Fft8::Tables *tab = lat_tables(st, &e);
    if (!tab) return (int)e;
    if (a.batch <= a.lat_cluster_max) {
        if (launch_lat2<LAT_BL, false>(a, tab, a.n, (size_t)0, st) == 0) return 0;
        (void)cudaGetLastError();      // a cluster that cannot be scheduled here (e.g. a partitioned device): one CTA per ciphertext instead
    }
    const size_t smem = lat_smem_bytes(a.n);
    auto kern = blind_rotate_lat_kernel<LAT_BL, false>;
    if ((e = ensure_dynamic_smem((const void *)kern, smem)) != cudaSuccess) return (int)e;
    kern<<<a.batch, LAT_THREADS, smem, st>>>((const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx,
                                             (const cplx *)a.bsk, (uint64_t *)a.out, a.out_stride, a.n, a.mode, tab, a.n, (size_t)0);
    return (int)cudaGetLastError();
}

// a.bsk: [batch][n_iter] summed Fourier GGSWs (key_ct_stride complex values per ciphertext); n_iter external products each
int br_lat_plain_launch(const Params &p, const BRLaunch &a, int n_iter, size_t key_ct_stride, cudaStream_t st) {
    if (!br_lat_plain_supported(p)) return -1;
    if (a.batch <= 0) return 0;
    cudaError_t e = cudaSuccess;
    const Fft8::Tables *tab = lat_tables(st, &e);
    if (!tab) return (int)e;
    const size_t smem = lat_smem_bytes(a.n);
#define X(BL)                                                                                                                       \
    if (p.pbs_base_log == BL && a.batch <= a.lat_cluster_max) {                                                                     \
        if (launch_lat2<BL, true>(a, tab, n_iter, key_ct_stride, st) == 0) return 0;                                                \
        (void)cudaGetLastError();                                                                                                   \
    }                                                                                                                               \
    if (p.pbs_base_log == BL) {                                                                                                     \
        auto kern = blind_rotate_lat_kernel<BL, true>;                                                                              \
        if ((e = ensure_dynamic_smem((const void *)kern, smem)) != cudaSuccess) return (int)e;                                      \
        kern<<<a.batch, LAT_THREADS, smem, st>>>((const uint64_t *)a.lwe_in, a.lwe_stride, (const uint64_t *)a.luts, a.lut_idx,     \
                                                 (const cplx *)a.bsk, (uint64_t *)a.out, a.out_stride, a.n, a.mode, tab, n_iter,   \
                                                 key_ct_stride);                                                                               \
        return (int)cudaGetLastError();                                                                                             \
    }
    LAT_PLAIN_INSTANCES(X)
#undef X
    return -1;
}

}  // namespace b200tfhe

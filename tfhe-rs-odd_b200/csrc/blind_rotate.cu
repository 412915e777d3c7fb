// blind_rotate.cu -- batched blind rotation / programmable bootstrap kernels for sm_100a.
//
// One launch per ciphertext batch.  A CTA owns CTS ciphertexts for all n CMUX iterations: their
// accumulators live in shared memory as integer torus polynomials, one team of R threads per GLWE
// polynomial.  Stage sequence per iteration (br_core.cuh): rotate-subtract + signed decomposition ->
// register/shared-memory forward FFT -> Fourier multiply-accumulate against this iteration's GGSW
// (read straight from L2: every CTA of a wave walks i in step) -> inverse FFT -> torus rounding.
// Modulus switch (prologue), LUT rotation by -b_hat (prologue) and sample extraction (epilogue) are fused.
//
// Replaces FourierLweBootstrapKeyView::{blind_rotate_assign, bootstrap, bootstrap_without_sample_extract}
// (tfhe/src/core_crypto/fft_impl/fft64/crypto/bootstrap.rs:243-412) and
// par_convert_standard_lwe_bootstrap_key_to_fourier (algorithms/lwe_bootstrap_key_conversion.rs:99-152).
#include "br_core.cuh"
#include "kernels.h"
#include <cstdlib>

#ifndef B200TFHE_REALIGN
#define B200TFHE_REALIGN 1
#endif

namespace b200tfhe {

template <typename T, int R, int K, int L, int BL, int CTS> struct BRConfig {
    typedef BRCore<T, R, K, L, BL> C;
    static constexpr int K1 = K + 1;
    static constexpr int THREADS = CTS * K1 * R;
    static_assert(THREADS % 32 == 0, "teams must fill whole warps");
    static constexpr size_t FBUF_BYTES = (size_t)K1 * L * C::REGION * sizeof(cplx);   // per ciphertext
    static constexpr size_t ACC_BYTES = (size_t)K1 * C::N * sizeof(T);
    static constexpr size_t TW_BYTES = sizeof(TwTables<R>);   // both twiddle tables, staged once per CTA
    static size_t smem_bytes(int n) {
        size_t ahat = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
        return CTS * (FBUF_BYTES + ACC_BYTES + ahat) + TW_BYTES;
    }
};

template <typename T, int R, int K, int L, int BL, int CTS>
__global__ void __launch_bounds__(CTS *(K + 1) * R)
blind_rotate_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                    const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                    size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                    const cplx *__restrict__ tw_inv_, size_t bsk_ct_stride, int bsk_ct_div) {
    typedef BRConfig<T, R, K, L, BL, CTS> CFG;
    typedef typename CFG::C C;
    constexpr int K1 = K + 1, N = C::N, REGION = C::REGION;
    constexpr bool REALIGN = B200TFHE_REALIGN;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int team = threadIdx.x / R, lane = threadIdx.x % R;
    const int ct_local = team / K1, r = team % K1;
    int ct = blockIdx.x * CTS + ct_local;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;  // ghost ciphertext: same work, no output (keeps barriers uniform)

    const size_t ahat_bytes = (((size_t)(n + 1) * sizeof(uint16_t)) + 15) & ~(size_t)15;
    cplx *fbuf_all = reinterpret_cast<cplx *>(smem_raw);
    T *acc_all = reinterpret_cast<T *>(smem_raw + CTS * CFG::FBUF_BYTES);
    uint16_t *ahat_all = reinterpret_cast<uint16_t *>(smem_raw + CTS * (CFG::FBUF_BYTES + CFG::ACC_BYTES));
    cplx *fbuf_ct = fbuf_all + (size_t)ct_local * K1 * L * REGION;
    cplx *fbuf_r = fbuf_ct + (size_t)r * L * REGION;
    T *acc_ct = acc_all + (size_t)ct_local * K1 * N;
    T *acc_r = acc_ct + (size_t)r * N;
    uint16_t *ahat = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(ahat_all) + ct_local * ahat_bytes);

    // twiddle tables: global -> shared once per CTA (short, fixed LDS latency instead of L1/L2 round trips)
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + CTS * (CFG::FBUF_BYTES + CFG::ACC_BYTES + ahat_bytes));
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += CFG::THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    // ---- prologue: modulus switch of the whole input LWE (fft_impl/common.rs:26-43) --------------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    bsk += (size_t)(ct / bsk_ct_div) * bsk_ct_stride;   // per-ciphertext GGSW list (0: the shared bootstrap key)
    for (int i = r * R + lane; i <= n; i += K1 * R) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    __syncthreads();
    // acc <- LUT * X^{-b_hat}  (polynomial_algorithms.rs:315-354), as a rotation by 2N - b_hat
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int c = lane; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncwarp();

    // ciphertext-wide barrier: only the K+1 teams of one ciphertext exchange data, so when they fill whole warps a
    // named barrier per ciphertext lets the CTS ciphertexts of the CTA drift apart instead of waiting on each other
    auto ct_barrier = [&]() {
#if defined(__CUDA_ARCH__)
        if constexpr (CTS > 1 && (K1 * R) % 32 == 0 && CTS <= 15) {
            asm volatile("bar.sync %0, %1;" ::"r"(ct_local + 1), "n"(K1 * R) : "memory");
        } else {
            __syncthreads();
        }
#endif
    };

    // ---- n CMUX iterations -----------------------------------------------------------------------------
    for (int i = 0; i < n; i++) {
        if constexpr (CTS > 1 && REALIGN) __syncthreads();   // re-align the CTA's ciphertexts (instruction-cache sharing)
        const int a = ahat[i];
        C::digits_fwd_pass1(lane, acc_r, a, fbuf_r, twf);
        __syncwarp();
        // bootstrap-key row 0 is fetched in parts across the second FFT pass and the barrier (L2 latency hiding)
        const cplx *ggsw_i = bsk + (size_t)i * C::GGSW_CPLX;
        cplx F[R];
        C::template mac_prefetch_part<0>(lane, r, ggsw_i, F);
#pragma unroll 1
        for (int t = 0; t < L; t++) {   // rolled: one copy of the second pass in the instruction stream
            cplx x[R];
            C::fwd_pass2_load(lane, x, fbuf_r + t * REGION);
            __syncwarp();
            C::fwd_pass2_store(lane, x, fbuf_r + t * REGION);
        }
        if constexpr (R > C::CH) C::template mac_prefetch_part<1>(lane, r, ggsw_i, F);
        ct_barrier();
        C::mac_finish(lane, r, ggsw_i, fbuf_ct, F);
        TeamFFT<R>::inv_pass1_compute(lane, F, twi);   // butterflies and middle twiddle BEFORE the barrier: only the store needs the region free
        ct_barrier();
        TeamFFT<R>::pass1_store(lane, F, fbuf_r);
        __syncwarp();
        C::inv_pass2_accumulate(lane, fbuf_r, acc_r);
        __syncwarp();
    }

    // ---- epilogue ---------------------------------------------------------------------------------------
    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = lane; c < N; c += R) o[(size_t)r * N + c] = acc_r[c];
    } else {
        // sample extract, coefficient 0 (algorithms/glwe_sample_extraction.rs:91-147)
        if (r < K) {
            for (int t = lane; t < N; t += R) o[(size_t)r * N + t] = (t == 0) ? acc_r[0] : (T)(0 - acc_r[N - t]);
        } else if (lane == 0) {
            o[(size_t)K * N] = acc_r[0];
        }
    }
}

// Standard-domain bootstrap key [n][l][k+1][k+1][N] -> device Fourier layout [n][l][k+1][k+1][p][lane]
template <typename T, int R>
__global__ void __launch_bounds__(R * (64 / R) * (R == 32 ? 1 : 2)) bsk_convert_kernel(const T *__restrict__ std_bsk, cplx *__restrict__ out,
                                                           size_t npoly, const cplx *__restrict__ tw_fwd_) {
    typedef BRCore<T, R, 1, 1, 1> C;
    constexpr int TEAMS = (R == 32) ? 2 : 8;  // 64 or 128 threads; keeps static smem < 48 KB
    __shared__ __align__(16) cplx regions[TEAMS][C::REGION];
    const int team = threadIdx.x / R, lane = threadIdx.x % R;
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_fwd_);
    size_t per_cta = (npoly + gridDim.x - 1) / gridDim.x;
    per_cta = (per_cta + TEAMS - 1) / TEAMS * TEAMS;
    size_t begin = (size_t)blockIdx.x * per_cta;
    for (size_t base = begin; base < begin + per_cta; base += TEAMS) {
        size_t pi = base + team;
        bool ok = pi < npoly;
        size_t pc = ok ? pi : npoly - 1;
        C::key_fwd_pass1(lane, std_bsk + pc * C::N, regions[team], twf);
        __syncwarp();
        cplx x[R];
        C::fwd_pass2_load(lane, x, regions[team]);
        __syncwarp();
        TeamFFT<R>::fwd_pass2_compute(x);
        if (ok) {
            cplx *o = out + pc * C::M;
            static_for<0, R>([&](auto p) { o[decltype(p)::value * R + lane] = x[decltype(p)::value]; });
        }
        __syncwarp();
    }
}

// ---- launch wrappers ----------------------------------------------------------------------------------------
template <typename T, int R, int K, int L, int BL, int CTS>
static int launch_br(const BRLaunch &a, cudaStream_t st) {
    typedef BRConfig<T, R, K, L, BL, CTS> CFG;
    auto kern = blind_rotate_kernel<T, R, K, L, BL, CTS>;
    size_t smem = CFG::smem_bytes(a.n);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    int grid = (a.batch + CTS - 1) / CTS;
    kern<<<grid, CFG::THREADS, smem, st>>>((const T *)a.lwe_in, a.lwe_stride, (const T *)a.luts, a.lut_idx,
                                           (const cplx *)a.bsk, (T *)a.out, a.out_stride, a.n, a.batch, a.mode,
                                           (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv, a.bsk_ct_stride, a.bsk_ct_div);
    return (int)cudaGetLastError();
}

// first two: WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS (shortint/parameters/parameters_wopbs_message_carry.rs:432-451; also 1_CARRY_1), its
// bootstrap (level 2, base 2^15) and the GGSWs its circuit bootstrap produces (cbs level 3, base 2^5); next two: the same pair for
// the (k 3, N 512) sets WOPBS_PARAM_MESSAGE_2_CARRY_1 / 4_CARRY_0 (pbs level 2 base 2^12, cbs level 4 base 2^3);
// last two: 1_CARRY_2 / 3_CARRY_1 / 4_CARRY_1 / 6_CARRY_0 (pbs level 3 base 2^9, cbs level 5 base 2^3)
#define BR_INSTANCES(X)              \
    X(uint64_t, 32, 1, 2, 15, 2)     \
    X(uint64_t, 32, 1, 3, 5, 1)      \
    X(uint64_t, 16, 3, 2, 12, 1)     \
    X(uint64_t, 16, 3, 4, 3, 1)      \
    X(uint64_t, 16, 3, 3, 9, 1)      \
    X(uint64_t, 16, 3, 5, 3, 1)      \
    X(uint64_t, 32, 1, 1, 23, 3)     \
    X(uint32_t, 16, 2, 3, 6, 2)      \
    X(uint64_t, 16, 3, 1, 18, 1)     \
    X(uint64_t, 16, 1, 2, 12, 1)

int br_supported(const Params &p) {
#define X(T, R, K, L, BL, CTS) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return R;
    BR_INSTANCES(X)
#undef X
    return br_big_supported(p);
}

// small batches (fewer ciphertexts than 3 per SM): one ciphertext per CTA spreads them over more SMs and avoids ghost
// ciphertexts inside a CTA -- this is also the single-PBS latency configuration
#define BR_SMALL_BATCH_INSTANCES(X)  \
    X(uint64_t, 32, 1, 1, 23, 1)

int br_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
    if (a.batch <= 0) return 0;
    // Measured on one B200 at PARAM_MESSAGE_2_CARRY_2 (tools/sweep_small_batches.sh): up to two ciphertexts per SM the
    // one-ciphertext-per-CTA latency kernel wins (148: 2.8 ms, 296: 5.6 ms); from there one wave of the TMEM kernel
    // (4 ciphertexts per CTA, 7.4 ms for anything up to 592) beats both it and the generic kernel.
    const bool per_ct_key = a.bsk_ct_stride != 0;   // only the generic kernel indexes the key per ciphertext
    if (per_ct_key || a.force == BR_GENERIC) {
#define X(T, R, K, L, BL, CTS) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) \
        return launch_br<T, R, K, L, BL, CTS>(a, st);
        BR_INSTANCES(X)
#undef X
        return per_ct_key ? -1 : br_big_launch(p, a, st);
    }
    // explicit choices (context option "br_kernel": parity tests exercise every kernel on the same inputs); -2: not built for this set
    if (a.force == BR_TMEM) return br_tmem_supported(p) ? br_tmem_launch(p, a, st) : -2;
    if (a.force == BR_LAT) return br_lat_supported(p) ? br_lat_launch(p, a, st) : -2;
    if (a.force == BR_X16) return br_x16_supported(p) ? br_x16_launch(p, a, st) : -2;
    if (a.force == BR_PAIR) return br_pair_supported(p) ? br_pair_launch(p, a, st) : -2;
#ifdef B200TFHE_BUILD_WS
    if (a.force == BR_WS) return br_ws_supported(p) ? br_ws_launch(p, a, st) : -2;
#else
    if (a.force == BR_WS) return -2;
#endif
    if (a.force == BR_AUTO) {
        if (a.batch <= a.lat_max && br_lat_supported(p)) return br_lat_launch(p, a, st);
        if (br_tmem_supported(p)) {
            // The TMEM kernel runs in waves of 4 x (number of SMs) ciphertexts and a partial last wave costs a full wave time (7.4 ms).
            // A remainder of up to lat_max ciphertexts is cheaper on the one-ciphertext-per-SM latency kernel (148: 2.8 ms, 296: 5.5 ms):
            // what a strong-scaling split of one batch over several GPUs leaves behind (2048 per GPU = 3 full waves + 272).
            int sms = 148, dev = 0;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            const int wave = 4 * sms, rem = a.batch % wave;
            if (rem > 0 && rem <= a.lat_max && a.batch > wave && br_lat_supported(p)) {
                BRLaunch main = a, tail = a;
                main.batch = a.batch - rem;
                const size_t esz = (size_t)p.bits / 8;
                tail.batch = rem;
                tail.lwe_in = (const char *)a.lwe_in + (size_t)main.batch * a.lwe_stride * esz;
                tail.out = (char *)a.out + (size_t)main.batch * a.out_stride * esz;
                if (a.lut_idx) tail.lut_idx = a.lut_idx + main.batch;
                if (int rc = br_tmem_launch(p, main, st)) return rc;
                a.launches = 2;
                return br_lat_launch(p, tail, st);
            }
            return br_tmem_launch(p, a, st);
        }
        if (br_x16_supported(p)) {
            // boolean shape: blind_rotate_x16b_kernel (8 ciphertexts per CTA, key through a TMA ring) runs a wave of 8 x SMs ciphertexts in
            // 11.75 ms, the generic kernel a wave of 4 x SMs in 7.85 ms (profiles/r02_x16_variants.json): take the one with less wave time
            int sms = 148, dev = 0;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            const int wx = (a.batch + 8 * sms - 1) / (8 * sms), wg = (a.batch + 4 * sms - 1) / (4 * sms);
            if (3 * wx <= 2 * wg) return br_x16_launch(p, a, st);
        }
        // PARAM_MESSAGE_1_CARRY_1 shape: six ciphertexts per CTA once there is more than a ciphertext per SM-slot of the generic kernel
        if (br_pair_supported(p) && a.batch >= 5 * 148) return br_pair_launch(p, a, st);
    }
    if (a.batch < 3 * 148 || a.force == BR_GENERIC_1CT) {
#define X(T, R, K, L, BL, CTS) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) \
        return launch_br<T, R, K, L, BL, CTS>(a, st);
        BR_SMALL_BATCH_INSTANCES(X)
#undef X
    }
#define X(T, R, K, L, BL, CTS) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) \
        return launch_br<T, R, K, L, BL, CTS>(a, st);
    BR_INSTANCES(X)
#undef X
    return br_big_launch(p, a, st);
}

int bsk_convert_polys_launch(int R, int bits, size_t npoly, const void *std_bsk, void *out, const void *tw_fwd, cudaStream_t st) {
    int grid = 148 * 4;
    if (R == 32 && bits == 64) bsk_convert_kernel<uint64_t, 32><<<grid, 64, 0, st>>>((const uint64_t *)std_bsk, (cplx *)out, npoly, (const cplx *)tw_fwd);
    else if (R == 16 && bits == 64) bsk_convert_kernel<uint64_t, 16><<<grid, 128, 0, st>>>((const uint64_t *)std_bsk, (cplx *)out, npoly, (const cplx *)tw_fwd);
    else if (R == 32 && bits == 32) bsk_convert_kernel<uint32_t, 32><<<grid, 64, 0, st>>>((const uint32_t *)std_bsk, (cplx *)out, npoly, (const cplx *)tw_fwd);
    else if (R == 16 && bits == 32) bsk_convert_kernel<uint32_t, 16><<<grid, 128, 0, st>>>((const uint32_t *)std_bsk, (cplx *)out, npoly, (const cplx *)tw_fwd);
    else return -1;
    return (int)cudaGetLastError();
}

int bsk_convert_launch(const Params &p, const void *std_bsk, void *out, const void *tw_fwd, cudaStream_t st) {
    size_t npoly = (size_t)p.n * p.pbs_level * (p.k + 1) * (p.k + 1);
    int R = br_supported(p);
    if (R == 256) return bsk_convert_big_launch(p, std_bsk, out, tw_fwd, st);
    return bsk_convert_polys_launch(R, p.bits, npoly, std_bsk, out, tw_fwd, st);
}

}  // namespace b200tfhe

// br_core.cuh -- the per-lane stages of one CMUX iteration of the blind rotation.
//
// Restates (as a different algorithm) fft64/crypto/bootstrap.rs:243-344 (blind_rotate_assign) and
// fft64/crypto/ggsw.rs:477-598 (add_external_product_assign) of the reference:
//   ct1 = acc * X^{a_hat} - acc          polynomial_algorithms.rs:425-490   (fused into the digit stage)
//   digits = signed decomposition(ct1)   fft64/math/decomposition.rs:27-86
//   D[r][lvl] = forward FFT(digits)      fft/mod.rs:496-515
//   F[j] = sum_{lvl desc, r asc} GGSW[lvl][r][j] (.) D[r][lvl]      ggsw.rs:616-697 (same summation order)
//   acc[j] += backward-as-torus(F[j])    fft/mod.rs:539-557
//
// A "team" of R threads owns one GLWE polynomial of one ciphertext; every function takes the lane index
// so that the CUDA kernel (blind_rotate.cu) and the CPU lane emulator (tests/native/emul.cpp) run the
// very same code.  Shared-memory objects per ciphertext:
//   T    acc[K+1][N]                 the accumulator (integer torus), the only state carried across iterations
//   cplx fbuf[K+1][L][REGION]        exchange buffer of each forward FFT, then its Fourier output D[r][lvl]
#pragma once
#include "fft_reg.cuh"

namespace b200tfhe {

// value of (P * X^a)[c] for a in [0, 2N): negacyclic rotation read
template <typename T, int N> HD T rot_read(const T *poly, int c, int a) {
    // s = c - a lies in (-2N, N); its low log2(2N) bits in two's complement are (c - a) mod 2N, so the index is
    // s mod N and the sign flips when bit log2(N) of s is set (X^N = -1)
    const int s = c - a;
    T v = poly[s & (N - 1)];
    return (s & N) ? (T)(0 - v) : v;
}

template <typename T, int R, int K, int L, int BASE_LOG> struct BRCore {
    typedef TeamFFT<R> FFT;
    static constexpr int M = R * R, N = 2 * M, K1 = K + 1, LOGN = clog2(N);
    static constexpr int REGION = FFT::REGION;
    static constexpr int GGSW_CPLX = L * K1 * K1 * M;  // one Fourier GGSW, in cplx
    typedef typename TorusTraits<T>::Signed S;

    // One-level decomposition in closed form (same values as decomp_init_state + decomp_next_digit, checked
    // exhaustively around the rounding / carry boundaries by tests/test_host_cpu.py via the emulator): the state
    // is the top BASE_LOG bits rounded to nearest, the digit is that state re-centred into (-B/2, B/2].
    HD static int32_t single_level_digit(T x) {
        constexpr int W = TorusTraits<T>::BITS;
        constexpr uint32_t MASK = (1u << BASE_LOG) - 1, HALF = 1u << (BASE_LOG - 1);
        const uint32_t top = (uint32_t)(x >> (W - BASE_LOG - 1));   // BASE_LOG + 1 bits (BASE_LOG <= 30)
        const uint32_t res = ((top + 1) >> 1) & MASK;
        return (int32_t)res - (int32_t)((res > HALF) ? (MASK + 1) : 0u);
    }

    // Stage A1: for polynomial `acc_r`, rotation a (in [0,2N)): compute all L digit planes of
    // acc*X^a - acc at this lane's R folded points, run forward pass 1 for each level and leave the
    // transposed partial results in fbuf_r[t] (t = 0 is level L, the first one the recurrence emits).
    HD static void digits_fwd_pass1(int lane, const T *acc_r, int a, cplx *fbuf_r, const cplx (*twf)[R]) {
        if constexpr (L == 1) {
            int32_t dre[L][R], dim[L][R];
            digits(lane, acc_r, a, dre, dim);
            cplx x[R];
            digits_to_cplx<0>(dre, dim, x);
            FFT::fwd_pass1(lane, x, fbuf_r, twf);
        } else {
            // several levels: the decomposition states of this lane's 2R coefficients stay in registers and ONE rolled loop
            // emits a level's digits and transforms them, so the instruction stream holds a single copy of the radix-R FFT
            // (the unrolled form missed the instruction cache: profiles/r01_ncu_other_kernels.json, boolean: no_instruction 1.5/issue)
            // a state holds L*BASE_LOG bits (+ a carry): 32-bit registers whenever that fits, whatever the torus width
            typedef typename StateWord<T, L * BASE_LOG>::type S;
            S s0[R], s1[R];
            static_for<0, R>([&](auto p) {
                constexpr int m = brev(decltype(p)::value, FFT::LOGR);
                const int c = lane + R * m;
                s0[decltype(p)::value] = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c, a) - acc_r[c], BASE_LOG, L);
                s1[decltype(p)::value] = (S)decomp_init_state<T>(rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M], BASE_LOG, L);
            });
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
            for (int t = 0; t < L; t++) {
                cplx x[R];
                static_for<0, R>([&](auto p) {
                    x[decltype(p)::value] = cplx{int_to_double((int32_t)decomp_next_digit<S>(s0[decltype(p)::value], BASE_LOG)),
                                                 int_to_double((int32_t)decomp_next_digit<S>(s1[decltype(p)::value], BASE_LOG))};
                });
                FFT::fwd_pass1(lane, x, fbuf_r + t * REGION, twf);
            }
        }
    }
    HD static void digits(int lane, const T *acc_r, int a, int32_t (&dre)[L][R], int32_t (&dim)[L][R]) {
        digits_range<0, R>(lane, acc_r, a, dre, dim);
    }
    template <int P0, int P1>
    HD static void digits_range(int lane, const T *acc_r, int a, int32_t (&dre)[L][R], int32_t (&dim)[L][R]) {
        static_for<P0, P1>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
            const int c = lane + R * m;
            T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c];
            T d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
            if constexpr (L == 1) {
                dre[0][decltype(p)::value] = single_level_digit(d0);
                dim[0][decltype(p)::value] = single_level_digit(d1);
            } else {
                T s0 = decomp_init_state<T>(d0, BASE_LOG, L), s1 = decomp_init_state<T>(d1, BASE_LOG, L);
                static_for<0, L>([&](auto t) {
                    dre[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s0, BASE_LOG);
                    dim[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s1, BASE_LOG);
                });
            }
        });
    }
    template <int TL> HD static void digits_to_cplx(const int32_t (&dre)[L][R], const int32_t (&dim)[L][R], cplx (&x)[R]) {
        static_for<0, R>([&](auto p) {
            x[decltype(p)::value] = cplx{int_to_double(dre[TL][decltype(p)::value]), int_to_double(dim[TL][decltype(p)::value])};
        });
    }

    // Stage A2 (after a team sync), one level at a time: load the transposed data ...
    HD static void fwd_pass2_load(int lane, cplx (&x)[R], const cplx *region) { FFT::fwd_pass2_load(lane, x, region); }
    // ... (team sync) ... second radix-R pass and store D in position order [p][lane] (unpadded, coalesced)
    HD static void fwd_pass2_store(int lane, cplx (&x)[R], cplx *region) {
        FFT::fwd_pass2_compute(x);
        static_for<0, R>([&](auto p) { region[decltype(p)::value * R + lane] = x[decltype(p)::value]; });
    }

    // Stage B1: Fourier multiply-accumulate for output polynomial j, in two calls so that the loads of the
    // bootstrap key (L2 latency, the dominant stall of the first version of this kernel: profiles/r01_*) are in
    // flight while the team waits at the ciphertext-wide barrier.
    // ggsw: this iteration's Fourier GGSW in device layout [lvl 0..L-1 = level 1..L][r][j][p][lane].
    // fbuf: the ciphertext's D buffers [r][t][...] with t = 0 <-> level L.
    // Rows are consumed in the reference's order (level L down to 1, r ascending: ggsw.rs:524-577).
    static constexpr int NROWS = L * K1;
    static constexpr int CH = 16;                 // positions per key chunk kept in registers
    HD static const cplx *ggsw_row(const cplx *ggsw, int q, int j) {
        const int t = q / K1, r = q % K1, lvl = L - 1 - t;
        return ggsw + ((size_t)(lvl * K1 + r) * K1 + j) * M;
    }
    // (a) BEFORE the barrier: fetch key row 0 into F (register pp <-> position brev(pp), what the DIT wants), in
    // R/CH parts so that the first part can be issued while the forward FFT's second pass still owns most registers
    template <int PART> HD static void mac_prefetch_part(int lane, int j, const cplx *ggsw, cplx (&F)[R]) {
        const cplx *g = ggsw_row(ggsw, 0, j);
        static_for<PART * CH, PART * CH + CH>([&](auto pp) {
            constexpr int pos = brev(decltype(pp)::value, FFT::LOGR);
            F[decltype(pp)::value] = ldg_stream(g + pos * R + lane);
        });
    }
    HD static void mac_prefetch(int lane, int j, const cplx *ggsw, cplx (&F)[R]) {
        static_for<0, R / CH>([&](auto part) { mac_prefetch_part<decltype(part)::value>(lane, j, ggsw, F); });
    }
    // (b) AFTER the barrier: F = G_0 (.) D_0 in place, then F += G_q (.) D_q chunk by chunk; a chunk is CH
    // positions of one key row, its loads are issued together (memory-level parallelism CH per thread) and the
    // first one is issued before row 0 is consumed so that its latency overlaps that arithmetic.
    static constexpr int HPR = R / CH;            // chunks per row
    static constexpr int NCHUNK = NROWS * HPR;
    template <int C_> HD static void load_chunk(int lane, int j, const cplx *ggsw, cplx (&h)[CH]) {
        constexpr int q = C_ / HPR, half = C_ % HPR;
        const cplx *g = ggsw_row(ggsw, q, j);
        static_for<0, CH>([&](auto i) {
            constexpr int pp = half * CH + decltype(i)::value;
            h[decltype(i)::value] = ldg_stream(g + brev(pp, FFT::LOGR) * R + lane);
        });
    }
    template <int C_> HD static void consume_chunk(int lane, const cplx *fbuf, cplx (&F)[R], const cplx (&h)[CH]) {
        constexpr int q = C_ / HPR, half = C_ % HPR;
        const cplx *d = fbuf + (size_t)((q % K1) * L + q / K1) * REGION;
        static_for<0, CH>([&](auto i) {
            constexpr int pp = half * CH + decltype(i)::value;
            mac_one(F[pp], h[decltype(i)::value], d[brev(pp, FFT::LOGR) * R + lane]);
        });
    }
    HD static void mac_finish(int lane, int j, const cplx *ggsw, const cplx *fbuf, cplx (&F)[R]) {
        cplx h[CH];
        if constexpr (NCHUNK > HPR) load_chunk<HPR>(lane, j, ggsw, h);
        static_for<0, R>([&](auto pp) {   // row 0 = (t 0, r 0): first Fourier buffer of the ciphertext
            constexpr int pos = brev(decltype(pp)::value, FFT::LOGR);
            F[decltype(pp)::value] = cmul(F[decltype(pp)::value], fbuf[pos * R + lane]);
        });
        static_for<HPR, NCHUNK>([&](auto c) {
            if constexpr (decltype(c)::value != HPR) load_chunk<decltype(c)::value>(lane, j, ggsw, h);
            consume_chunk<decltype(c)::value>(lane, fbuf, F, h);
        });
    }
    HD static void mac_one(cplx &f, cplx gv, cplx dv) {
        f.re = fma_(gv.re, dv.re, f.re);
        f.re = fma_(-gv.im, dv.im, f.re);
        f.im = fma_(gv.re, dv.im, f.im);
        f.im = fma_(gv.im, dv.re, f.im);
    }

    // ---- pieces used by the TMEM variant of the kernel (blind_rotate.cu, single level): the team's own forward
    // transform is still in registers after pass 2, so its own key row is applied register-to-register; the other
    // team's transform arrives through tensor memory.  Register pp <-> Fourier position brev(pp) throughout.
    template <int HALF> HD static void own_chunk_load(int lane, int j, const cplx *ggsw, cplx (&g)[CH]) {
        const cplx *row = ggsw + ((size_t)j * K1 + j) * M;          // level 1, row r = j, column j
        static_for<0, CH>([&](auto i) {
            constexpr int pp = HALF * CH + decltype(i)::value;
            g[decltype(i)::value] = ldg_stream(row + brev(pp, FFT::LOGR) * R + lane);
        });
    }
    // F[pp] = D_own[pp] * g   (xb holds D_own already permuted to register order pp)
    template <int HALF> HD static void own_chunk_apply(const cplx (&xb)[R], const cplx (&g)[CH], cplx (&F)[R]) {
        static_for<0, CH>([&](auto i) {
            constexpr int pp = HALF * CH + decltype(i)::value;
            F[pp] = cmul(xb[pp], g[decltype(i)::value]);
        });
    }
    // quarter QT (8 positions, register order pp = QT*8 + e <-> Fourier position brev(pp)) of key row (r, j), level 1
    template <int QT> HD static void quarter_load(int lane, int r, int j, const cplx *ggsw, cplx (&h)[8]) {
        const cplx *row = ggsw + ((size_t)r * K1 + j) * M;
        static_for<0, 8>([&](auto e) {
            constexpr int pp = QT * 8 + decltype(e)::value;
            h[decltype(e)::value] = ldg_stream(row + brev(pp, FFT::LOGR) * R + lane);
        });
    }
#if defined(__CUDA_ARCH__)
    // L1 prefetch of that quarter: 8 position rows x 512 B = 32 lines, one per lane
    template <int QT> DEV static void quarter_prefetch(int lane, int r, int j, const cplx *ggsw) {
        const cplx *row = ggsw + ((size_t)r * K1 + j) * M;
        const int pp = QT * 8 + (lane >> 2);
        const int pos = (int)(__brev((unsigned)pp) >> (32 - FFT::LOGR));
        prefetch_l1(reinterpret_cast<const char *>(row + pos * R) + (lane & 3) * 128);
    }
    // L1 prefetch of the 8 KB chunk HALF of key row (r, j): 64 lines, two per lane
    template <int HALF> DEV static void chunk_prefetch(int lane, int r, int j, const cplx *ggsw) {
        const cplx *row = ggsw + ((size_t)r * K1 + j) * M;
        static_for<0, CH / 8>([&](auto h) {
            const int pp = HALF * CH + decltype(h)::value * 8 + (lane >> 2);
            const int pos = (int)(__brev((unsigned)pp) >> (32 - FFT::LOGR));
            prefetch_l1(reinterpret_cast<const char *>(row + pos * R) + (lane & 3) * 128);
        });
    }
#endif
    template <int HALF> HD static void other_chunk_load(int lane, int r, int j, const cplx *ggsw, cplx (&h)[CH]) {
        const cplx *row = ggsw + ((size_t)r * K1 + j) * M;
        static_for<0, CH>([&](auto i) {
            constexpr int pp = HALF * CH + decltype(i)::value;
            h[decltype(i)::value] = ldg_stream(row + brev(pp, FFT::LOGR) * R + lane);
        });
    }

    // Stage B2 (after a ciphertext-wide sync: every team finished reading D): inverse pass 1 into own region
    HD static void inv_pass1(int lane, cplx (&F)[R], cplx *region, const cplx (*twi)[R]) { FFT::inv_pass1(lane, F, region, twi); }

    // Stage B3 (after a team sync): inverse pass 2, round to the torus, accumulate
    HD static void inv_pass2_accumulate(int lane, const cplx *region, T *acc_j) {
        cplx x[R];
        FFT::inv_pass2(lane, x, region);
        static_for<0, R>([&](auto m) {
            const int c = lane + R * decltype(m)::value;
            acc_j[c] += torus_from_double<T>(x[decltype(m)::value].re);
            acc_j[c + M] += torus_from_double<T>(x[decltype(m)::value].im);
        });
    }

    // Forward transform of one standard-domain key polynomial (forward_as_torus, fft/mod.rs:197-218),
    // pass 1.  Used by the bootstrap-key conversion kernel.
    HD static void key_fwd_pass1(int lane, const T *poly, cplx *region, const cplx (*twf)[R]) {
        cplx x[R];
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
            const int c = lane + R * m;
            x[decltype(p)::value] = cplx{torus_to_double<T>(poly[c]), torus_to_double<T>(poly[c + M])};
        });
        FFT::fwd_pass1(lane, x, region, twf);
    }
};

// signed decomposition of the polynomial d_r as it is (no rotate-and-subtract) + first forward pass, level by level with the
// decomposition states of this lane's 2R coefficients in registers (the rolled form of BRCore::digits_fwd_pass1)
template <typename T, int R, int K, int L, int BL>
HD void plain_digits_fwd_pass1(int lane, const T *d_r, cplx *fbuf_r, const cplx (*twf)[R]) {
    typedef BRCore<T, R, K, L, BL> C;
    typedef typename C::FFT FFT;
    typedef typename StateWord<T, L * BL>::type S;
    S s0[R], s1[R];
    static_for<0, R>([&](auto p) {
        constexpr int m = brev(decltype(p)::value, FFT::LOGR);
        const int c = lane + R * m;
        s0[decltype(p)::value] = (S)decomp_init_state<T>(d_r[c], BL, L);
        s1[decltype(p)::value] = (S)decomp_init_state<T>(d_r[c + C::M], BL, L);
    });
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int t = 0; t < L; t++) {
        cplx x[R];
        static_for<0, R>([&](auto p) {
            x[decltype(p)::value] = cplx{int_to_double((int32_t)decomp_next_digit<S>(s0[decltype(p)::value], BL)),
                                         int_to_double((int32_t)decomp_next_digit<S>(s1[decltype(p)::value], BL))};
        });
        FFT::fwd_pass1(lane, x, fbuf_r + t * C::REGION, twf);
    }
}

}  // namespace b200tfhe

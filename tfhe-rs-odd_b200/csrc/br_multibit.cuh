// br_multibit.cuh -- the per-lane stages of one iteration of the MULTI-BIT blind rotation.
//
// Restates (as a different algorithm) algorithms/lwe_multi_bit_programmable_bootstrapping.rs of the reference:
//   prepare_multi_bit_ggsw_mem_optimized :18-85     GGSW_sum = GGSW_0 + sum_{idx>=1} GGSW_idx (.) FFT(X^{d_idx}),
//                                                   d_idx = modswitch(sum of the mask elements selected by idx)
//   multi_bit_deterministic_blind_rotate_assign     acc <- external_product(GGSW_sum, acc), one per group of G mask
//   :548-...  (and the threaded variant :295-546)   elements: n/G sequential iterations instead of n
// A bootstrap key holds 2^G GGSW ciphertexts per group (lwe_multi_bit_bootstrap_key_generation.rs:87-172,401-427).
//
// What is done differently here: GGSW_sum is never materialised.  By linearity
//     F_j = sum_{lvl,r} GGSW_sum[lvl][r][j] (.) D[r][lvl] = sum_idx  zeta^{d_idx} (.) ( sum_{lvl,r} GGSW_idx[lvl][r][j] (.) D[r][lvl] )
// so every thread runs the ordinary Fourier multiply-accumulate against each of the 2^G key ciphertexts on a chunk of its
// frequencies and combines the partial sums with the monomial's Fourier coefficients.  Those are powers of the evaluation
// points of the negacyclic transform: frequency f = lane + R*pos evaluates at zeta_f = e^{i pi (1 - 4f)/N}, hence
//     FFT(X^d)[f] = zeta_f^d = e^{i pi d (1 - 4 lane)/N} * e^{-2 pi i (d*pos mod R)/R}            (N = 2 R^2)
// = one per-thread base root per (iteration, idx) times an entry of a table of the R-th roots of unity.
#pragma once
#include "br_core.cuh"
#include <math.h>

namespace b200tfhe {

// e^{i pi e/N}, e integer (any sign), N a power of two: the argument e/N is exact in f64
HD cplx unit_root_pi(int e, int N) {
#if defined(__CUDA_ARCH__)
    double s, c;
    sincospi((double)e / (double)N, &s, &c);
    return cplx{c, s};
#else
    const long double a = 3.14159265358979323846264338327950288L * (long double)e / (long double)N;
    return cplx{(double)cosl(a), (double)sinl(a)};
#endif
}

template <typename T, int R, int K, int L, int BASE_LOG, int G> struct MBCore : BRCore<T, R, K, L, BASE_LOG> {
    typedef BRCore<T, R, K, L, BASE_LOG> C;
    typedef typename C::FFT FFT;
    static constexpr int M = C::M, N = C::N, K1 = C::K1, REGION = C::REGION, LOGN = C::LOGN, NROWS = C::NROWS;
    static constexpr int NG = 1 << G;         // GGSW ciphertexts per group (ggsw_per_multi_bit_element)
    static constexpr int NMONO = NG - 1;      // monomials per group (idx 0 is the constant term)
    static constexpr int CH = 8;              // frequencies per chunk kept in registers
    static constexpr size_t GROUP_CPLX = (size_t)NG * C::GGSW_CPLX;

    // monomial degrees of one group: deg[idx-1] = modswitch(sum_b bit_{G-1-b}(idx) * a_b) mod 2N   (:47-64)
    HD static uint16_t group_degree(const T *mask_elems, int idx) {
        T s = 0;
        for (int b = 0; b < G; b++)
            if ((idx >> (G - (b + 1))) & 1) s += mask_elems[b];
        return (uint16_t)(modswitch<T>(s, LOGN) & (2 * N - 1));
    }

    // Stage A1: signed decomposition of acc itself (the multi-bit loop has no rotate-and-subtract) + forward pass 1
    HD static void digits_plain_fwd_pass1(int lane, const T *acc_r, cplx *fbuf_r, const cplx (*twf)[R]) {
        int32_t dre[L][R], dim[L][R];
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
            const int c = lane + R * m;
            const T d0 = acc_r[c], d1 = acc_r[c + M];
            if constexpr (L == 1) {
                dre[0][decltype(p)::value] = C::single_level_digit(d0);
                dim[0][decltype(p)::value] = C::single_level_digit(d1);
            } else {
                T s0 = decomp_init_state<T>(d0, BASE_LOG, L), s1 = decomp_init_state<T>(d1, BASE_LOG, L);
                static_for<0, L>([&](auto t) {
                    dre[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s0, BASE_LOG);
                    dim[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s1, BASE_LOG);
                });
            }
        });
        static_for<0, L>([&](auto t) {
            cplx x[R];
            C::template digits_to_cplx<decltype(t)::value>(dre, dim, x);
            FFT::fwd_pass1(lane, x, fbuf_r + decltype(t)::value * REGION, twf);
        });
    }

    // table of the R-th roots of unity used by the monomial factors: wr[m] = e^{-2 pi i m/R}
    HD static cplx wr_entry(int m) { return unit_root_pi(-4 * m * R, N); }   // pi * (-4 m R)/N = -2 pi m/R  since N = 2 R^2

    // Stage B1: F (register pp <-> Fourier position brev(pp)) for output polynomial j.
    //   group: the NG Fourier GGSWs of this iteration, each in device layout [lvl][r][j][pos][lane]
    //   fbuf:  the ciphertext's D buffers [r][t] (t = 0 <-> level L), position order, as written by fwd_pass2_store
    //   deg:   the NMONO monomial degrees of this group
    HD static void mb_mac(int lane, int j, const cplx *group, const cplx *fbuf, const uint16_t *deg, const cplx *wr, cplx (&F)[R]) {
        // idx = 0: the constant term, no monomial factor
        static_for<0, R / CH>([&](auto ck) {
            cplx t[CH];
            partial_mac<decltype(ck)::value>(lane, j, group, fbuf, t);
            static_for<0, CH>([&](auto i) { F[decltype(ck)::value * CH + decltype(i)::value] = t[decltype(i)::value]; });
        });
        // idx >= 1.  G = 3: a rolled loop (one copy of the MAC code in the instruction cache: the unrolled form ran at
        // no_instruction 1.7 / issue); G = 2: unrolled (measured 49.3k vs 45.3k KS+PBS/s rolled).
        // An L1 prefetch of the next chunk (one line per lane) was measured and rejected: 27.7k / 36.4k vs 49.3k / 41.7k.
        if constexpr (G <= 2) {
            cplx base[NMONO];
            int dgs[NMONO];
            static_for<0, NMONO>([&](auto g) {
                dgs[decltype(g)::value] = deg[decltype(g)::value];
                base[decltype(g)::value] = unit_root_pi((dgs[decltype(g)::value] * (1 - 4 * lane)) & (2 * N - 1), N);
            });
            static_for<0, R / CH>([&](auto ck) {
                static_for<1, NG>([&](auto g) {
                    cplx t[CH];
                    partial_mac<decltype(ck)::value>(lane, j, group + (size_t) decltype(g)::value * C::GGSW_CPLX, fbuf, t);
                    static_for<0, CH>([&](auto i) {
                        constexpr int pp = decltype(ck)::value * CH + decltype(i)::value;
                        constexpr int pos = brev(pp, FFT::LOGR);
                        const cplx w = wr[(dgs[decltype(g)::value - 1] * pos) & (R - 1)];
                        C::mac_one(F[pp], cmul(base[decltype(g)::value - 1], w), t[decltype(i)::value]);
                    });
                });
            });
        } else {
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
            for (int g = 1; g < NG; g++) {
                const int dg = deg[g - 1];
                const cplx base = unit_root_pi((dg * (1 - 4 * lane)) & (2 * N - 1), N);
                const cplx *ggsw = group + (size_t)g * C::GGSW_CPLX;
                static_for<0, R / CH>([&](auto ck) {
                    cplx t[CH];
                    partial_mac<decltype(ck)::value>(lane, j, ggsw, fbuf, t);
                    static_for<0, CH>([&](auto i) {
                        constexpr int pp = decltype(ck)::value * CH + decltype(i)::value;
                        constexpr int pos = brev(pp, FFT::LOGR);
                        const cplx w = wr[(dg * pos) & (R - 1)];
                        C::mac_one(F[pp], cmul(base, w), t[decltype(i)::value]);
                    });
                });
            }
        }
    }
    // t[i] = sum_{lvl desc, r asc} GGSW[lvl][r][j][pos_i] * D[r][lvl][pos_i] for the CH positions of chunk CK
    template <int CK> HD static void partial_mac(int lane, int j, const cplx *ggsw, const cplx *fbuf, cplx (&t)[CH]) {
        static_for<0, NROWS>([&](auto q) {
            const cplx *row = C::ggsw_row(ggsw, decltype(q)::value, j);
            const cplx *d = fbuf + (size_t)((decltype(q)::value % K1) * L + decltype(q)::value / K1) * REGION;
            static_for<0, CH>([&](auto i) {
                constexpr int pos = brev(CK * CH + decltype(i)::value, FFT::LOGR);
                const cplx h = ldg_stream(row + pos * R + lane);
                const cplx dv = d[pos * R + lane];
                if constexpr (decltype(q)::value == 0) t[decltype(i)::value] = cmul(h, dv);
                else C::mac_one(t[decltype(i)::value], h, dv);
            });
        });
    }

    // Stage B3: inverse pass 2, round to the torus, STORE (add_external_product_assign into a zero-filled destination)
    HD static void inv_pass2_store(int lane, const cplx *region, T *acc_j) {
        cplx x[R];
        FFT::inv_pass2(lane, x, region);
        static_for<0, R>([&](auto m) {
            const int c = lane + R * decltype(m)::value;
            acc_j[c] = torus_from_double<T>(x[decltype(m)::value].re);
            acc_j[c + M] = torus_from_double<T>(x[decltype(m)::value].im);
        });
    }
};

}  // namespace b200tfhe

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
    int s = c - a;
    s += (s < 0) ? 2 * N : 0;
    T v = poly[s & (N - 1)];
    return (s & N) ? (T)(0 - v) : v;
}

template <typename T, int R, int K, int L, int BASE_LOG> struct BRCore {
    typedef TeamFFT<R> FFT;
    static constexpr int M = R * R, N = 2 * M, K1 = K + 1, LOGN = clog2(N);
    static constexpr int REGION = FFT::REGION;
    static constexpr int GGSW_CPLX = L * K1 * K1 * M;  // one Fourier GGSW, in cplx
    typedef typename TorusTraits<T>::Signed S;

    // Stage A1: for polynomial `acc_r`, rotation a (in [0,2N)): compute all L digit planes of
    // acc*X^a - acc at this lane's R folded points, run forward pass 1 for each level and leave the
    // transposed partial results in fbuf_r[t] (t = 0 is level L, the first one the recurrence emits).
    HD static void digits_fwd_pass1(int lane, const T *acc_r, int a, cplx *fbuf_r, const cplx (*twf)[R]) {
        int32_t dre[L][R], dim[L][R];
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, FFT::LOGR);
            const int c = lane + R * m;
            T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c];
            T d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
            T s0 = decomp_init_state<T>(d0, BASE_LOG, L), s1 = decomp_init_state<T>(d1, BASE_LOG, L);
            static_for<0, L>([&](auto t) {
                dre[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s0, BASE_LOG);
                dim[decltype(t)::value][decltype(p)::value] = (int32_t)decomp_next_digit<T>(s1, BASE_LOG);
            });
        });
        static_for<0, L>([&](auto t) {
            cplx x[R];
            static_for<0, R>([&](auto p) {
                x[decltype(p)::value] = cplx{int_to_double(dre[decltype(t)::value][decltype(p)::value]),
                                             int_to_double(dim[decltype(t)::value][decltype(p)::value])};
            });
            FFT::fwd_pass1(lane, x, fbuf_r + decltype(t)::value * REGION, twf);
        });
    }

    // Stage A2 (after a team sync), one level at a time: load the transposed data ...
    HD static void fwd_pass2_load(int lane, cplx (&x)[R], const cplx *region) { FFT::fwd_pass2_load(lane, x, region); }
    // ... (team sync) ... second radix-R pass and store D in position order [p][lane] (unpadded, coalesced)
    HD static void fwd_pass2_store(int lane, cplx (&x)[R], cplx *region) {
        FFT::fwd_pass2_compute(x);
        static_for<0, R>([&](auto p) { region[decltype(p)::value * R + lane] = x[decltype(p)::value]; });
    }

    // Stage B1 (after a ciphertext-wide sync): Fourier multiply-accumulate for output polynomial j.
    // ggsw: this iteration's Fourier GGSW in device layout [lvl 0..L-1 = level 1..L][r][j][p][lane].
    // fbuf: the ciphertext's D buffers [r][t][...] with t = 0 <-> level L.
    HD static void mac(int lane, int j, const cplx *ggsw, const cplx *fbuf, cplx (&F)[R]) {
        static_for<0, R>([&](auto p) { F[decltype(p)::value] = cplx{0.0, 0.0}; });
        for (int t = 0; t < L; t++) {          // level L down to 1, like ggsw.into_levels().rev()
            const int lvl = L - 1 - t;
            for (int r = 0; r < K1; r++) {
                const cplx *g = ggsw + ((size_t)(lvl * K1 + r) * K1 + j) * M;
                const cplx *d = fbuf + (size_t)(r * L + t) * REGION;
                static_for<0, R>([&](auto pp) {
                    // DIT wants register pp to hold position brev(pp)
                    constexpr int pos = brev(decltype(pp)::value, FFT::LOGR);
                    cplx gv = g[pos * R + lane], dv = d[pos * R + lane];
                    cplx &f = F[decltype(pp)::value];
                    f.re = fma_(gv.re, dv.re, f.re);
                    f.re = fma_(-gv.im, dv.im, f.re);
                    f.im = fma_(gv.re, dv.im, f.im);
                    f.im = fma_(gv.im, dv.re, f.im);
                });
            }
        }
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

}  // namespace b200tfhe

// fft3.cuh -- three-pass complex FFT of size 1024 for a team of 64 threads holding 16 points each.
//
// The single-ciphertext (latency) blind rotation spreads one polynomial over 64 threads instead of 32 so that one
// ciphertext occupies all four schedulers of an SM; with 16 points per thread 1024 = 16 x 4 x 16 needs three
// register passes and two shared-memory exchanges.  Same mathematics as fft_reg.cuh (fold, twist e^{i pi j/N}, DFT kernel
// e^{-2 pi i jf/M}; reference: fft_impl/fft64/math/fft/mod.rs:197-326,496-557) and the same Fourier-domain order as the
// bootstrap key built by bsk_convert_kernel: value f of the transform is "position" f.
//
// Index algebra (w_n = e^{-2 pi i/n}); input index j = j1 + 16 (u + 4 v), output index f = 64 f1 + 16 g1 + g2:
//   pass A  thread tau = j1 + 16 u :  Z[tau][g2]      = sum_v  x[tau + 64 v] w_16^{v g2}          (radix 16)
//                                     Z'              = Z * ta[g2][tau]          ta = (pre-twist of tau) * w_64^{u g2}
//   pass B  thread beta = j1 + 16 gq: Y[j1][16g1+g2]  = sum_u  Z'[j1 + 16 u][g2] w_4^{u g1}       (radix 4, g2 = 4 gq + e)
//                                     Y'              = Y * tb[4e+g1][beta]      tb = w_1024^{j1 f2}, f2 = 16 g1 + g2
//   pass C  thread t = f2:            X[64 f1 + f2]   = sum_j1 Y'[j1][f2] w_16^{j1 f1}            (radix 16)
// The inverse runs the SAME forward code on conjugated data (inv(X) = conj(fwd(conj X)) / M); its normalisation and
// the part of the inverse twist that depends on the thread are folded into tb[1].
//
// Everything is __host__ __device__ so that tests/native/emul.cpp can execute it thread by thread.
#pragma once
#include "fft_reg.cuh"

namespace b200tfhe {

struct Fft3 {
    static constexpr int M = 1024, N = 2048, P = 16, T = 64;
    static constexpr int ROW = 17;            // padded row of an exchange buffer (cplx): conflict-free 16-byte accesses
    static constexpr int BUF = T * ROW;       // cplx per exchange buffer

    struct Tables {
        cplx ta_fwd[16][64];   // [g2][tau]        e^{i pi tau/N} w_64^{(tau>>4) g2}
        cplx ta_inv[16][4];    // [g2][u]          w_64^{u g2}
        cplx tb[2][16][64];    // [dir][4e+g1][beta]   w_1024^{j1 f2};  dir 1: times e^{i pi f2/N} / M
    };

    // Forward only: multiply the folded input by the part of the twist that depends on the register (e^{i pi 64 v/N}).
    // x[p] holds element v = brev4(p) of this thread, i.e. coefficient tau + 64 v (+ M as imaginary part).
    HD static void pre_twist(cplx (&x)[16]) {
        static_for<0, 16>([&](auto p) { x[decltype(p)::value] = cmul_root<2 * brev(decltype(p)::value, 4)>(x[decltype(p)::value]); });
    }
    // Inverse only: x[q] <- x[q] e^{i pi 64 q/N} (the data is the conjugate of the result, hence the sign).
    HD static void post_twist(cplx (&x)[16]) {
        static_for<0, 16>([&](auto q) { x[decltype(q)::value] = cmul_root<2 * decltype(q)::value>(x[decltype(q)::value]); });
    }

    // ta: table base; entry g2 of this thread is ta[g2 * ta_stride + ta_idx] (forward: stride 64, idx tau; inverse: 4, tau>>4)
    HD static void pass_a(int tau, cplx (&x)[16], const cplx *ta, int ta_stride, int ta_idx, cplx *buf1) {
        cplx w[16];   // fetched before the butterflies so that their latency hides behind the arithmetic
        static_for<0, 16>([&](auto g2) { w[decltype(g2)::value] = ta[decltype(g2)::value * ta_stride + ta_idx]; });
        fft_dit<16, -1>(x);
        static_for<0, 16>([&](auto g2) {
            constexpr int g = decltype(g2)::value;
            buf1[tau * ROW + g] = cmul(x[g], w[g]);
        });
    }
    // team sync before and after
    HD static void pass_b(int beta, const cplx *tb, const cplx *buf1, cplx *buf2) {
        const int j1 = beta & 15, gq = beta >> 4;
        cplx y[4][4], w[4][4];
        static_for<0, 16>([&](auto i) {
            constexpr int e = decltype(i)::value >> 2, pu = decltype(i)::value & 3;
            y[e][pu] = buf1[(j1 + 16 * brev(pu, 2)) * ROW + 4 * gq + e];
        });
        static_for<0, 16>([&](auto i) { w[decltype(i)::value >> 2][decltype(i)::value & 3] = tb[decltype(i)::value * 64 + beta]; });
        static_for<0, 4>([&](auto e_) {
            constexpr int e = decltype(e_)::value;
            fft_dit<4, -1>(y[e]);
            static_for<0, 4>([&](auto g1) {
                constexpr int g = decltype(g1)::value;
                buf2[(16 * g + 4 * gq + e) * ROW + j1] = cmul(y[e][g], w[e][g]);
            });
        });
    }
    // x[f1] = X[64 f1 + t]
    HD static void pass_c(int t, cplx (&x)[16], const cplx *buf2) {
        static_for<0, 16>([&](auto p) { x[decltype(p)::value] = buf2[t * ROW + brev(decltype(p)::value, 4)]; });
        fft_dit<16, -1>(x);
    }
};

}  // namespace b200tfhe

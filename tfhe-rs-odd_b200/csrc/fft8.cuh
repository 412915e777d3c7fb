// fft8.cuh -- three-pass complex FFT of size 1024 for a team of 128 threads holding 8 points each.
//
// The single-ciphertext (latency) blind rotation gives one polynomial to FOUR warps, so that an SM working on one ciphertext
// (two polynomials) has two warps per scheduler, one of each polynomial: while one waits for a shared-memory exchange the
// other issues butterflies.  (Round 1 used 64 threads x 16 points, one warp per scheduler: the shared-memory data pipe and the
// FP64 pipe then work strictly one after the other, profiles/r01_ncu_latency.json.)  With 8 points per thread
// 1024 = 8 x 16 x 8: radix 8 in registers, radix 16 as radix 8 in registers + one radix-2 stage between two neighbouring
// threads through warp shuffles, radix 8 in registers; two shared-memory exchanges.  Same mathematics as fft_reg.cuh (fold, twist
// e^{i pi j/N}, DFT kernel e^{-2 pi i jf/M}; reference: fft_impl/fft64/math/fft/mod.rs:197-326,496-557) and the same Fourier-domain
// order as the bootstrap key built by bsk_convert_kernel: value f of the transform is "position" f.
//
// Index algebra (w_n = e^{-2 pi i/n}); input index j = tau + 128 v, tau = tau0 + 8 u; output index f = g + 8 h + 128 f2:
//   pass A  thread tau:                  Z[tau][g]    = sum_v  x[tau + 128 v] w_8^{v g}                         (radix 8)
//                                        Z'           = Z * ta[g][tau]             ta = (pre-twist of tau) * w_128^{u g}    (factored, see Tables)
//   pass B  thread beta = (tau0,half,g): E|O[q]       = sum_m  Z'[tau0 + 8 (2m + half)][g] w_8^{m q}            (radix 8)
//                                        Y[tau0][g][h] = E[q] +- w_16^q O[q],  h = q | q + 8   (radix 2 across the thread pair:
//                                                        half 0 ends up with q = 0..3, half 1 with q = 4..7, both signs each)
//                                        Y'           = Y * tb[slot][beta]         tb = w_1024^{tau0 c},  c = g + 8 h       (factored, see TwB)
//   pass C  thread c:                    X[c + 128 f2] = sum_tau0 Y'[tau0][c] w_8^{tau0 f2}                      (radix 8)
// The output layout (thread c holds X[c + 128 f2]) is the input layout (thread tau holds x[tau + 128 v]), so the Fourier
// product feeds the inverse transform without another exchange.  The inverse runs the SAME forward code on conjugated data
// (inv(X) = conj(fwd(conj X)) / M); its normalisation and the part of the inverse twist that depends on the thread are folded
// into tb[1].
//
// Everything is __host__ __device__ so that tests/native/emul.cpp can execute it thread by thread (the shuffle is the
// caller's: pass_b_first leaves the four values to send, pass_b_second takes the four received ones).
#pragma once
#include "fft_reg.cuh"

namespace b200tfhe {

struct Fft8 {
    static constexpr int M = 1024, N = 2048, P = 8, T = 128;
    static constexpr int BUF1 = 8 * T;        // cplx, first exchange buffer [g][tau]: 16-byte accesses of 8 consecutive lanes hit all banks
    static constexpr int ROW2 = 9;            // padded row of the second exchange buffer [c][tau0] (cplx)
    static constexpr int BUF2 = T * ROW2;

    // The twiddles are kept factored so that a thread can hold all it needs in registers: the shared-memory data pipe, not the
    // FP64 pipe, bounds the latency kernel, and full per-thread tables cost 16 sixteen-byte loads per thread and transform.
    //   pass A: ta[g] = twist(tau) * w_128^{u g}: one per-thread factor (forward only) times a value shared by the 8 threads of a u
    //   pass B: tb[slot] = w_1024^{tau0 c}, c = c0 + 8 i + 64 j (c0 = g + 32 half, i = slot & 3, j = slot >> 2)
    //                    = A * B_i * C^j,  A = w_1024^{tau0 c0}, B_i = w_128^{tau0 i}, C = w_16^{tau0}
    //           inverse: times e^{i pi c/N} / M, split the same way into A, B_i and C (each factor one correctly rounded value)
    struct TwB { cplx a, b1, b2, b3, c; };
    struct Tables {
        cplx twist[128];       // [tau]           e^{i pi tau/N}
        cplx ta[8][16];        // [g][u]          w_128^{u g}
        TwB tb[2][128];        // [dir][beta]
    };

    // which h the thread pair member `half` owns in slot s (s < 4: the "+" output of q = 4 half + s, s >= 4: its "-" output)
    HD static constexpr int slot_h(int half, int s) { return 4 * half + (s & 3) + 8 * (s >> 2); }

    // Forward only: multiply the folded input by the part of the twist that depends on the register (e^{i pi 128 v/N}).
    // x[p] holds element v = brev3(p) of this thread, i.e. coefficient tau + 128 v (+ M as imaginary part).
    HD static void pre_twist(cplx (&x)[8]) {
        static_for<0, 8>([&](auto p) { x[decltype(p)::value] = cmul_root<4 * brev(decltype(p)::value, 3)>(x[decltype(p)::value]); });
    }
    // Inverse only: x[q] <- x[q] e^{i pi 128 q/N} (the data is the conjugate of the result, hence the sign).
    HD static void post_twist(cplx (&x)[8]) {
        static_for<0, 8>([&](auto q) { x[decltype(q)::value] = cmul_root<4 * decltype(q)::value>(x[decltype(q)::value]); });
    }

    // ta: the [g][u] table; twist: this thread's e^{i pi tau/N} (forward) -- the inverse passes with_twist = false
    HD static void pass_a(int tau, cplx (&x)[8], const cplx (*ta)[16], cplx twist, bool with_twist, cplx *buf1) {
        cplx w[8];   // fetched before the butterflies so that their latency hides behind the arithmetic
        static_for<1, 8>([&](auto g) { w[decltype(g)::value] = ta[decltype(g)::value][tau >> 3]; });
        fft_dit<8, -1>(x);
        if (with_twist) {
            buf1[tau] = cmul(x[0], twist);
            static_for<1, 8>([&](auto g) { buf1[decltype(g)::value * T + tau] = cmul(x[decltype(g)::value], cmul(twist, w[decltype(g)::value])); });
        } else {
            buf1[tau] = x[0];
            static_for<1, 8>([&](auto g) { buf1[decltype(g)::value * T + tau] = cmul(x[decltype(g)::value], w[decltype(g)::value]); });
        }
    }
    // team sync before.  z: this thread's half of a 16-point transform; send: what the pair partner (beta ^ 8) needs
    HD static void pass_b_first(int beta, const cplx *buf1, cplx (&z)[8], cplx (&send)[4]) {
        const int tau0 = beta & 7, half = (beta >> 3) & 1, g = beta >> 4;
        static_for<0, 8>([&](auto p) { z[decltype(p)::value] = buf1[g * T + tau0 + 8 * (2 * brev(decltype(p)::value, 3) + half)]; });
        fft_dit<8, -1>(z);
        if (half) {   // the odd half carries the radix-2 twiddle w_16^q
            static_for<1, 8>([&](auto q) { z[decltype(q)::value] = cmul_root<-8 * decltype(q)::value>(z[decltype(q)::value]); });
        }
        static_for<0, 4>([&](auto i) { send[decltype(i)::value] = half ? z[decltype(i)::value] : z[4 + decltype(i)::value]; });
    }
    // recv: the partner's `send`.  Team sync after.
    HD static void pass_b_second(int beta, const cplx (&z)[8], const cplx (&recv)[4], const TwB &tw, cplx *buf2) {
        const int tau0 = beta & 7, half = (beta >> 3) & 1, g = beta >> 4;
        static_for<0, 4>([&](auto i_) {
            constexpr int i = decltype(i_)::value;
            const cplx a = half ? recv[i] : z[i], b = half ? z[4 + i] : recv[i];       // a = E[q], b = w_16^q O[q], q = 4 half + i
            const cplx yp{a.re + b.re, a.im + b.im}, ym{a.re - b.re, a.im - b.im};
            const int c = g + 8 * (4 * half + i);
            cplx w = tw.a;
            if constexpr (i == 1) w = cmul(w, tw.b1);
            if constexpr (i == 2) w = cmul(w, tw.b2);
            if constexpr (i == 3) w = cmul(w, tw.b3);
            buf2[c * ROW2 + tau0] = cmul(yp, w);
            buf2[(c + 64) * ROW2 + tau0] = cmul(ym, cmul(w, tw.c));
        });
    }
    // x[f2] = X[c + 128 f2]
    HD static void pass_c(int c, cplx (&x)[8], const cplx *buf2) {
        static_for<0, 8>([&](auto p) { x[decltype(p)::value] = buf2[c * ROW2 + brev(decltype(p)::value, 3)]; });
        fft_dit<8, -1>(x);
    }
};

}  // namespace b200tfhe

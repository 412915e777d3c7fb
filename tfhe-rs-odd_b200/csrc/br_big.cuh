// br_big.cuh -- CMUX stages for N = 8192 (PARAM_MESSAGE_3_CARRY_3: k = 1, l = 2): one CTA of 256 threads per ciphertext.
//
// Same reference semantics as br_core.cuh (fft64/crypto/ggsw.rs:477-598, bootstrap.rs:243-344); different shape because
// the accumulator (128 KB) and a Fourier polynomial (64 KB) no longer fit a warp-sized team:
//   * FFT of size M = 4096 = 16 x 256: every thread holds 16 points; pass A is a radix-16 DIT over the stride-256
//     index in registers, then a CTA-wide transpose through one 64 KB shared buffer hands each 16-thread team one
//     256-point DFT, done as 16 x 16 exactly like the small-N team FFT (in-place XOR-swizzled 16x16 transpose).
//   * After the forward transform thread `tid` holds positions p*256 + tid (p < 16) of EVERY polynomial, so the Fourier
//     multiply-accumulate is register-to-register: F[j][p] += G[..][j][p][tid] * D[p]; no Fourier buffer in shared memory.
//   * Shared memory: acc 128 KB + exchange 64 KB + mod-switched mask.
#pragma once
#include "fft_reg.cuh"
#include "br_core.cuh"

namespace b200tfhe {

struct BigTables {
    // middle twiddle of the 16 x 256 split, u_t * w_4096^{t k1} (u_t = e^{i pi t/N}, w_4096 = e^{-2 pi i/4096}), factored as
    // ta[k1 & 3][t] * tq[k1 >> 2][t] so that both factors (32 KB) live in shared memory: the full 2 x 64 KB tables were read
    // from L2 on every transform (96 of 224 global loads per thread and iteration, profiles/r01_ncu_other_kernels.json).
    // The inverse uses the conjugate; its 1/M is folded into the digit conversion (an exact power of two).
    cplx ta[4][256];    // u_t * w_4096^{t b}
    cplx tq[4][256];    // w_4096^{4 t a}
    cplx pf[8][16];     // plain 256-point team DFT: [b<4][lane] = w_256^{lane b}, [4+a][lane] = w_256^{lane 4a}
    cplx pi[8][16];     // conjugates
};

template <typename T, int K, int L, int BASE_LOG> struct BigCore {
    static constexpr int P = 16, TH = 256, M = 4096, N = 8192, K1 = K + 1, LOGP = 4, LOGN = 13;
    static constexpr int TWQ = 2;  // e^{i pi 256 m / N} = e^{2 pi i (2m)/128}
    static constexpr int GGSW_CPLX = L * K1 * K1 * M;
    static constexpr size_t ACC_BYTES = (size_t)K1 * N * sizeof(T), XCH_BYTES = (size_t)M * sizeof(cplx);

    HD static int swz(int row, int col) { return row * 16 + (col ^ row); }

    // digits of level index t (t = 0 <-> level L, the first digit the recurrence emits) of acc_r*X^a - acc_r
    template <int TL> HD static void digits(int tid, const T *acc_r, int a, cplx (&x)[P]) {
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            const int c = tid + TH * m;
            T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c];
            T d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
            typedef typename StateWord<T, L * BASE_LOG>::type S;   // 32-bit recurrence when the state fits
            S s0 = (S)decomp_init_state<T>(d0, BASE_LOG, L), s1 = (S)decomp_init_state<T>(d1, BASE_LOG, L);
            int32_t g0 = 0, g1 = 0;
            static_for<0, TL + 1>([&](auto) {
                g0 = (int32_t)decomp_next_digit<S>(s0, BASE_LOG);
                g1 = (int32_t)decomp_next_digit<S>(s1, BASE_LOG);
            });
            x[decltype(p)::value] = cplx{int_to_double_scaled<12>(g0), int_to_double_scaled<12>(g1)};   // digits / M
        });
    }
    // same with the level index at run time (rolled level loop of the kernel: one copy of the forward transform)
    HD static void digits_rt(int tid, const T *acc_r, int a, int tl, cplx (&x)[P]) {
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            const int c = tid + TH * m;
            T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c];
            T d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
            typedef typename StateWord<T, L * BASE_LOG>::type S;
            S s0 = (S)decomp_init_state<T>(d0, BASE_LOG, L), s1 = (S)decomp_init_state<T>(d1, BASE_LOG, L);
            int32_t g0 = 0, g1 = 0;
            static_for<0, L>([&](auto e) {    // every level's digit is a few integer instructions: compute all, select one
                const int32_t e0 = (int32_t)decomp_next_digit<S>(s0, BASE_LOG), e1 = (int32_t)decomp_next_digit<S>(s1, BASE_LOG);
                if (tl == decltype(e)::value) { g0 = e0; g1 = e1; }
            });
            x[decltype(p)::value] = cplx{int_to_double_scaled<12>(g0), int_to_double_scaled<12>(g1)};   // digits / M
        });
    }
    // digits of level index tl of the polynomial src_r as it is (stand-alone external product / CMUX: no rotate-and-subtract)
    HD static void digits_plain_rt(int tid, const T *src_r, int tl, cplx (&x)[P]) {
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            const int c = tid + TH * m;
            typedef typename StateWord<T, L * BASE_LOG>::type S;
            S s0 = (S)decomp_init_state<T>(src_r[c], BASE_LOG, L), s1 = (S)decomp_init_state<T>(src_r[c + M], BASE_LOG, L);
            int32_t g0 = 0, g1 = 0;
            static_for<0, L>([&](auto e) {
                const int32_t e0 = (int32_t)decomp_next_digit<S>(s0, BASE_LOG), e1 = (int32_t)decomp_next_digit<S>(s1, BASE_LOG);
                if (tl == decltype(e)::value) { g0 = e0; g1 = e1; }
            });
            x[decltype(p)::value] = cplx{int_to_double_scaled<12>(g0), int_to_double_scaled<12>(g1)};   // digits / M
        });
    }
    // L = 2: both levels' digits from ONE pass over the coefficients; the first (level L) goes to x, the second is packed
    // (two 16-bit digits, |digit| <= 2^(BASE_LOG-1) <= 2^15 - 1 requires BASE_LOG <= 15) for the step that transforms it later
    HD static void digits_both(int tid, const T *acc_r, int a, cplx (&x)[P], uint32_t (&packed)[P]) {
        static_assert(L != 2 || BASE_LOG <= 15, "packed second-level digits are 16 bits wide");
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            const int c = tid + TH * m;
            T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c];
            T d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
            typedef typename StateWord<T, L * BASE_LOG>::type S;
            S s0 = (S)decomp_init_state<T>(d0, BASE_LOG, L), s1 = (S)decomp_init_state<T>(d1, BASE_LOG, L);
            const int32_t g0 = (int32_t)decomp_next_digit<S>(s0, BASE_LOG), g1 = (int32_t)decomp_next_digit<S>(s1, BASE_LOG);
            const int32_t h0 = (int32_t)decomp_next_digit<S>(s0, BASE_LOG), h1 = (int32_t)decomp_next_digit<S>(s1, BASE_LOG);
            x[decltype(p)::value] = cplx{int_to_double_scaled<12>(g0), int_to_double_scaled<12>(g1)};
            packed[decltype(p)::value] = ((uint32_t)h0 & 0xFFFFu) | ((uint32_t)h1 << 16);
        });
    }
    HD static void digits_unpack(const uint32_t (&packed)[P], cplx (&x)[P]) {
        static_for<0, P>([&](auto p) {
            const uint32_t w = packed[decltype(p)::value];
            x[decltype(p)::value] = cplx{int_to_double_scaled<12>((int32_t)(int16_t)(w & 0xFFFFu)), int_to_double_scaled<12>((int32_t)w >> 16)};
        });
    }
    HD static void key_input(int tid, const T *poly, cplx (&x)[P]) {
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            const int c = tid + TH * m;
            x[decltype(p)::value] = cplx{torus_to_double<T>(poly[c]), torus_to_double<T>(poly[c + M])};
        });
    }

    // ---- forward: fwdA | CTA sync | fwdB_load | team sync | fwdB_pass1 | team sync | fwdB_pass2 | (CTA sync before next fwdA)
    // ta / tq: the two factor tables ([4][256] each), in shared memory for the blind rotation
    template <int K1_> HD static cplx mid_tw(int tid, const cplx (*ta)[TH], const cplx (*tq)[TH]) {
        if constexpr ((K1_ >> 2) == 0) return ta[K1_ & 3][tid];
        else return cmul(ta[K1_ & 3][tid], tq[K1_ >> 2][tid]);
    }
    HD static void fwdA(int tid, cplx (&x)[P], cplx *S1, const cplx (*ta)[TH], const cplx (*tq)[TH]) {
        fwdA_compute(tid, x, ta, tq);
        fwdA_store(tid, x, S1);
    }
    // the two halves of fwdA: everything that stays in registers (twist, radix-16 pass, middle twiddle), and the exchange store --
    // a kernel can compute the first half of the NEXT transform while the exchange of the current one is in flight
    HD static void fwdA_compute(int tid, cplx (&x)[P], const cplx (*ta)[TH], const cplx (*tq)[TH]) {
        static_for<0, P>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGP);
            x[decltype(p)::value] = cmul_root<m * TWQ>(x[decltype(p)::value]);
        });
        fft_dit<P, -1>(x);
        static_for<0, P>([&](auto k1) { x[decltype(k1)::value] = cmul(x[decltype(k1)::value], mid_tw<decltype(k1)::value>(tid, ta, tq)); });
    }
    HD static void fwdA_store(int tid, const cplx (&x)[P], cplx *S1) {
        static_for<0, P>([&](auto k1) { S1[decltype(k1)::value * TH + tid] = x[decltype(k1)::value]; });
    }
    HD static void fwdB_load(int tid, cplx (&x)[P], const cplx *S1) {
        const int team = tid >> 4, lane = tid & 15;
        static_for<0, P>([&](auto p) { x[decltype(p)::value] = S1[team * TH + lane + 16 * brev(decltype(p)::value, LOGP)]; });
    }
    template <int DIR> HD static void team_twiddle(int lane, cplx (&x)[P], const cplx (*tw)[16]) {
        cplx r[4];
        for (int b = 1; b < 4; b++) r[b] = tw[b][lane];
        static_for<0, 4>([&](auto a) {
            if constexpr (decltype(a)::value == 0) {
                static_for<1, 4>([&](auto b) { x[decltype(b)::value] = cmul(x[decltype(b)::value], r[decltype(b)::value]); });
            } else {
                cplx q = tw[4 + decltype(a)::value][lane];
                x[decltype(a)::value * 4] = cmul(x[decltype(a)::value * 4], q);
                static_for<1, 4>([&](auto b) {
                    constexpr int e = decltype(a)::value * 4 + decltype(b)::value;
                    x[e] = cmul(x[e], cmul(r[decltype(b)::value], q));
                });
            }
        });
    }
    HD static void fwdB_pass1(int tid, cplx (&x)[P], cplx *S1, const BigTables *tb) {
        const int team = tid >> 4, lane = tid & 15;
        fft_dit<P, -1>(x);
        team_twiddle<-1>(lane, x, tb->pf);
        static_for<0, P>([&](auto e) { S1[team * TH + swz(decltype(e)::value, lane)] = x[decltype(e)::value]; });
    }
    // after this x[p] = Fourier position p*256 + tid
    HD static void fwdB_pass2(int tid, cplx (&x)[P], const cplx *S1) {
        const int team = tid >> 4, lane = tid & 15;
        static_for<0, P>([&](auto p) { x[decltype(p)::value] = S1[team * TH + swz(lane, brev(decltype(p)::value, LOGP))]; });
        fft_dit<P, -1>(x);
    }

    // F[j][p] += G[lvl][r][j][p*256 + tid] * x[p]
    HD static void mac(int tid, int lvl, int r, const cplx *ggsw, const cplx (&x)[P], cplx (&F)[K1][P]) {
        static_for<0, K1>([&](auto j) {
            const cplx *g = ggsw + ((size_t)(lvl * K1 + r) * K1 + decltype(j)::value) * M + tid;
            cplx gv[P];
            static_for<0, P>([&](auto p) { gv[decltype(p)::value] = ldg_stream(g + decltype(p)::value * TH); });
            static_for<0, P>([&](auto p) {
                cplx &f = F[decltype(j)::value][decltype(p)::value];
                const cplx a = gv[decltype(p)::value], d = x[decltype(p)::value];
                f.re = fma_(a.re, d.re, f.re);
                f.re = fma_(-a.im, d.im, f.re);
                f.im = fma_(a.re, d.im, f.im);
                f.im = fma_(a.im, d.re, f.im);
            });
        });
    }

    // f[p] += g[p] * x[p] for one output polynomial (same operation order as mac)
    HD static void mac_one_poly(const cplx (&g)[P], const cplx (&x)[P], cplx (&f)[P]) {
        static_for<0, P>([&](auto p) {
            const cplx a = g[decltype(p)::value], d = x[decltype(p)::value];
            cplx &o = f[decltype(p)::value];
            o.re = fma_(a.re, d.re, o.re);
            o.re = fma_(-a.im, d.im, o.re);
            o.im = fma_(a.re, d.im, o.im);
            o.im = fma_(a.im, d.re, o.im);
        });
    }

    // ---- inverse: (x[pp] = F[brev(pp)]) invA | team sync | invB_load | team sync | invB_store | CTA sync | invC
    HD static void invA(int tid, cplx (&x)[P], cplx *S1, const BigTables *tb) {
        invA_compute(tid, x, tb);
        invA_store(tid, x, S1);
    }
    // the two halves of invA: registers only, and the exchange store
    HD static void invA_compute(int tid, cplx (&x)[P], const BigTables *tb) {
        fft_dit<P, +1>(x);
        team_twiddle<+1>(tid & 15, x, tb->pi);
    }
    HD static void invA_store(int tid, const cplx (&x)[P], cplx *S1) {
        const int team = tid >> 4, lane = tid & 15;
        static_for<0, P>([&](auto l) { S1[team * TH + swz(decltype(l)::value, lane)] = x[decltype(l)::value]; });
    }
    HD static void invB_load(int tid, cplx (&x)[P], const cplx *S1) {
        const int team = tid >> 4, lane = tid & 15;
        static_for<0, P>([&](auto p) { x[decltype(p)::value] = S1[team * TH + swz(lane, brev(decltype(p)::value, LOGP))]; });
    }
    HD static void invB_store(int tid, cplx (&x)[P], cplx *S1) {
        const int team = tid >> 4, lane = tid & 15;
        fft_dit<P, +1>(x);
        static_for<0, P>([&](auto m2) { S1[team * TH + lane + 16 * decltype(m2)::value] = x[decltype(m2)::value]; });
    }
    // result: x[m].re -> coefficient tid + 256 m, x[m].im -> that + M, un-twisted (the 1/M was applied to the digits)
    HD static void invC(int tid, cplx (&x)[P], const cplx *S1, const cplx (*ta)[TH], const cplx (*tq)[TH]) {
        static_for<0, P>([&](auto pp) {
            constexpr int k1 = brev(decltype(pp)::value, LOGP);
            const cplx w = mid_tw<k1>(tid, ta, tq);
            x[decltype(pp)::value] = cmul(S1[k1 * TH + tid], cplx{w.re, -w.im});
        });
        fft_dit<P, +1>(x);
        static_for<0, P>([&](auto m) { x[decltype(m)::value] = cmul_root<-decltype(m)::value * TWQ>(x[decltype(m)::value]); });
    }
    HD static void accumulate(int tid, const cplx (&x)[P], T *acc_j) {
        static_for<0, P>([&](auto m) {
            const int c = tid + TH * decltype(m)::value;
            acc_j[c] += torus_from_double<T>(x[decltype(m)::value].re);
            acc_j[c + M] += torus_from_double<T>(x[decltype(m)::value].im);
        });
    }
};

}  // namespace b200tfhe

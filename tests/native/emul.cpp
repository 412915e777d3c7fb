// emul.cpp -- CPU lane emulator of the CUDA blind-rotation stages (TEST INFRASTRUCTURE).
// Compiles tfhe-rs-odd_b200/csrc/{fft_reg,br_core}.cuh as host code and runs the stage sequence of
// blind_rotate.cu with the R lanes of every team executed one after another; "syncs" are loop boundaries.
// Lets the no-GPU test-suite check the kernel's mathematics against the oracle.
#include "br_core.cuh"
#include "twiddles.hpp"
#include <vector>
#include <cstring>
#include <type_traits>
using namespace b200tfhe;

template <typename T, int R, int K, int L, int BL>
static void convert_bsk(const uint64_t *bsk_std, int n, std::vector<cplx> &out) {
    typedef BRCore<T, R, K, L, BL> C;
    static TwTables<R> tw; fill_tw_tables(tw);
    const int N = C::N, M = C::M, K1 = C::K1;
    size_t npoly = (size_t)n * L * K1 * K1;
    out.assign(npoly * M, cplx{0, 0});
    std::vector<cplx> region(C::REGION);
    std::vector<T> poly(N);
    cplx x[R][R];
    for (size_t pi = 0; pi < npoly; pi++) {
        for (int c = 0; c < N; c++) poly[c] = (T)bsk_std[pi * N + c];
        for (int lane = 0; lane < R; lane++) C::key_fwd_pass1(lane, poly.data(), region.data(), tw.fwd);
        for (int lane = 0; lane < R; lane++) C::fwd_pass2_load(lane, x[lane], region.data());
        for (int lane = 0; lane < R; lane++) {
            TeamFFT<R>::fwd_pass2_compute(x[lane]);
            for (int p = 0; p < R; p++) out[pi * M + p * R + lane] = x[lane][p];
        }
    }
}

template <typename T, int R, int K, int L, int BL>
static void blind_rotate(const std::vector<cplx> &fbsk, int n, const uint64_t *lwe, const uint64_t *lut, uint64_t *out_glwe) {
    typedef BRCore<T, R, K, L, BL> C;
    static TwTables<R> tw; fill_tw_tables(tw);
    const int N = C::N, K1 = C::K1;
    std::vector<T> acc((size_t)K1 * N), lutT((size_t)K1 * N);
    for (size_t c = 0; c < lutT.size(); c++) lutT[c] = (T)lut[c];
    int bhat = (int)(modswitch<T>((T)lwe[n], C::LOGN) & (2 * N - 1));
    int a0 = (2 * N - bhat) & (2 * N - 1);
    for (int r = 0; r < K1; r++)
        for (int c = 0; c < N; c++) acc[(size_t)r * N + c] = rot_read<T, N>(&lutT[(size_t)r * N], c, a0);
    std::vector<cplx> fbuf((size_t)K1 * L * C::REGION);
    static cplx x[R][R];
    static cplx F[K + 1][R][R];
    for (int i = 0; i < n; i++) {
        int a = (int)(modswitch<T>((T)lwe[i], C::LOGN) & (2 * N - 1));
        if (a == 0) continue;
        for (int r = 0; r < K1; r++)
            for (int lane = 0; lane < R; lane++)
                C::digits_fwd_pass1(lane, &acc[(size_t)r * N], a, &fbuf[(size_t)r * L * C::REGION], tw.fwd);
        for (int r = 0; r < K1; r++)
            for (int t = 0; t < L; t++) {
                cplx *reg = &fbuf[(size_t)(r * L + t) * C::REGION];
                for (int lane = 0; lane < R; lane++) C::fwd_pass2_load(lane, x[lane], reg);
                for (int lane = 0; lane < R; lane++) C::fwd_pass2_store(lane, x[lane], reg);
            }
        const cplx *ggsw = &fbsk[(size_t)i * C::GGSW_CPLX];
        for (int j = 0; j < K1; j++)
            for (int lane = 0; lane < R; lane++) { C::mac_prefetch(lane, j, ggsw, F[j][lane]); C::mac_finish(lane, j, ggsw, fbuf.data(), F[j][lane]); }
        for (int j = 0; j < K1; j++)
            for (int lane = 0; lane < R; lane++) C::inv_pass1(lane, F[j][lane], &fbuf[(size_t)j * L * C::REGION], tw.inv);
        for (int j = 0; j < K1; j++)
            for (int lane = 0; lane < R; lane++) C::inv_pass2_accumulate(lane, &fbuf[(size_t)j * L * C::REGION], &acc[(size_t)j * N]);
    }
    for (size_t c = 0; c < acc.size(); c++) out_glwe[c] = (uint64_t)acc[c];
}

template <typename T, int R, int K, int L, int BL>
static int run(const uint64_t *bsk_std, int n, const uint64_t *lwe, int batch, const uint64_t *lut, uint64_t *out) {
    std::vector<cplx> f;
    convert_bsk<T, R, K, L, BL>(bsk_std, n, f);
    const size_t glwe = (size_t)(K + 1) * 2 * R * R;
    for (int b = 0; b < batch; b++) blind_rotate<T, R, K, L, BL>(f, n, lwe + (size_t)b * (n + 1), lut, out + b * glwe);
    return 0;
}

// all arrays carried as uint64_t like the oracle's; bits selects the torus width
extern "C" int emul_blind_rotate(int n, int k, int N, int base_log, int level, int bits, const uint64_t *bsk_std,
                                 const uint64_t *lwe, int batch, const uint64_t *lut, uint64_t *out_glwe) {
#define CASE(T, R, K, L, BL) \
    if (bits == (int)sizeof(T) * 8 && N == 2 * R * R && k == K && level == L && base_log == BL) \
        return run<T, R, K, L, BL>(bsk_std, n, lwe, batch, lut, out_glwe);
    CASE(uint64_t, 16, 1, 2, 12)
    CASE(uint32_t, 16, 2, 3, 6)
    CASE(uint64_t, 32, 1, 1, 23)
    CASE(uint64_t, 16, 3, 1, 18)
#undef CASE
    return 1;
}

// forward FFT of one integer polynomial then backward-add into zero: roundtrip / product helpers for FFT unit tests
extern "C" int emul_fft_roundtrip(int N, const uint64_t *poly, uint64_t *out) {
    if (N == 2048) {
        typedef BRCore<uint64_t, 32, 1, 1, 23> C; constexpr int R = 32;
        static TwTables<R> tw; fill_tw_tables(tw);
        std::vector<cplx> region(C::REGION); std::vector<uint64_t> acc(N, 0);
        static cplx x[R][R];
        for (int lane = 0; lane < R; lane++) C::key_fwd_pass1(lane, poly, region.data(), tw.fwd);
        for (int lane = 0; lane < R; lane++) C::fwd_pass2_load(lane, x[lane], region.data());
        for (int lane = 0; lane < R; lane++) TeamFFT<R>::fwd_pass2_compute(x[lane]);
        // registers hold position p natural; inverse wants register pp = position brev(pp)
        static cplx y[R][R];
        for (int lane = 0; lane < R; lane++) for (int pp = 0; pp < R; pp++) y[lane][pp] = x[lane][brev(pp, 5)];
        for (int lane = 0; lane < R; lane++) C::inv_pass1(lane, y[lane], region.data(), tw.inv);
        for (int lane = 0; lane < R; lane++) C::inv_pass2_accumulate(lane, region.data(), acc.data());
        memcpy(out, acc.data(), sizeof(uint64_t) * N);
        return 0;
    }
    return 1;
}

// closed-form single-level digit vs the generic recurrence (returns the number of mismatches)
extern "C" int emul_check_single_level_digit(const uint64_t *xs, int count) {
    int bad = 0;
    for (int i = 0; i < count; i++) {
        uint64_t st = decomp_init_state<uint64_t>(xs[i], 23, 1);
        int32_t ref = (int32_t)decomp_next_digit<uint64_t>(st, 23);
        if (BRCore<uint64_t, 32, 1, 1, 23>::single_level_digit(xs[i]) != ref) bad++;
        uint64_t st2 = decomp_init_state<uint64_t>(xs[i], 18, 1);
        int32_t ref2 = (int32_t)decomp_next_digit<uint64_t>(st2, 18);
        if (BRCore<uint64_t, 16, 3, 1, 18>::single_level_digit(xs[i]) != ref2) bad++;
    }
    return bad;
}

// ---- N = 8192 (one 256-thread CTA per ciphertext): same stage sequence as blind_rotate_big.cu ----------------------
#include "br_big.cuh"
template <typename T, int K, int L, int BL>
static int run_big(const uint64_t *bsk_std, int n, const uint64_t *lwe, int batch, const uint64_t *lut, uint64_t *out) {
    typedef BigCore<T, K, L, BL> C;
    constexpr int K1 = K + 1, N = C::N, M = C::M, P = C::P, TH = 256;
    static BigTables tb; fill_big_tables(tb);
    std::vector<cplx> S1(M);
    static cplx x[TH][P];
    static cplx F[K1][TH][P];
    // key conversion
    size_t npoly = (size_t)n * L * K1 * K1;
    std::vector<cplx> fbsk(npoly * M);
    std::vector<T> poly(N);
    auto forward = [&]() {
        for (int t = 0; t < TH; t++) C::fwdA(t, x[t], S1.data(), tb.ta, tb.tq);
        for (int t = 0; t < TH; t++) C::fwdB_load(t, x[t], S1.data());
        for (int t = 0; t < TH; t++) C::fwdB_pass1(t, x[t], S1.data(), &tb);
        for (int t = 0; t < TH; t++) C::fwdB_pass2(t, x[t], S1.data());
    };
    for (size_t pi = 0; pi < npoly; pi++) {
        for (int c = 0; c < N; c++) poly[c] = (T)bsk_std[pi * N + c];
        for (int t = 0; t < TH; t++) C::key_input(t, poly.data(), x[t]);
        forward();
        for (int t = 0; t < TH; t++) for (int p = 0; p < P; p++) fbsk[pi * M + p * TH + t] = x[t][p];
    }
    std::vector<T> acc((size_t)K1 * N), lutT((size_t)K1 * N);
    for (size_t c = 0; c < lutT.size(); c++) lutT[c] = (T)lut[c];
    for (int b = 0; b < batch; b++) {
        const uint64_t *ct = lwe + (size_t)b * (n + 1);
        int bhat = (int)(modswitch<T>((T)ct[n], C::LOGN) & (2 * N - 1));
        int a0 = (2 * N - bhat) & (2 * N - 1);
        for (int r = 0; r < K1; r++)
            for (int c = 0; c < N; c++) acc[(size_t)r * N + c] = rot_read<T, N>(&lutT[(size_t)r * N], c, a0);
        for (int i = 0; i < n; i++) {
            int a = (int)(modswitch<T>((T)ct[i], C::LOGN) & (2 * N - 1));
            const cplx *ggsw = &fbsk[(size_t)i * C::GGSW_CPLX];
            for (int j = 0; j < K1; j++) for (int t = 0; t < TH; t++) for (int p = 0; p < P; p++) F[j][t][p] = cplx{0, 0};
            static_for<0, L>([&](auto tl) {
                for (int r = 0; r < K1; r++) {
                    for (int t = 0; t < TH; t++) C::template digits<decltype(tl)::value>(t, &acc[(size_t)r * N], a, x[t]);
                    forward();
                    for (int t = 0; t < TH; t++) {
                        cplx Ft[K1][P];
                        for (int j = 0; j < K1; j++) for (int p = 0; p < P; p++) Ft[j][p] = F[j][t][p];
                        C::mac(t, L - 1 - decltype(tl)::value, r, ggsw, x[t], Ft);
                        for (int j = 0; j < K1; j++) for (int p = 0; p < P; p++) F[j][t][p] = Ft[j][p];
                    }
                }
            });
            for (int j = 0; j < K1; j++) {
                for (int t = 0; t < TH; t++) for (int pp = 0; pp < P; pp++) x[t][pp] = F[j][t][brev(pp, C::LOGP)];
                for (int t = 0; t < TH; t++) C::invA(t, x[t], S1.data(), &tb);
                for (int t = 0; t < TH; t++) C::invB_load(t, x[t], S1.data());
                for (int t = 0; t < TH; t++) C::invB_store(t, x[t], S1.data());
                for (int t = 0; t < TH; t++) { C::invC(t, x[t], S1.data(), tb.ta, tb.tq); C::accumulate(t, x[t], &acc[(size_t)j * N]); }
            }
        }
        for (size_t c = 0; c < acc.size(); c++) out[(size_t)b * K1 * N + c] = (uint64_t)acc[c];
    }
    return 0;
}

extern "C" int emul_blind_rotate_big(int n, int k, int N, int base_log, int level, int bits, const uint64_t *bsk_std,
                                     const uint64_t *lwe, int batch, const uint64_t *lut, uint64_t *out_glwe) {
    if (bits == 64 && N == 8192 && k == 1 && level == 2 && base_log == 15)
        return run_big<uint64_t, 1, 2, 15>(bsk_std, n, lwe, batch, lut, out_glwe);
    return 1;
}

// ---- latency kernel (blind_rotate_lat.cu): 128 threads per polynomial, three-pass FFT of fft8.cuh ---------------------
#include "fft8.cuh"
// one 1024-point transform of the team: x[t][p] in pass-A register order on entry, pass-C order (x[t][f2] = X[t + 128 f2]) on exit
static void fft8_team(cplx (*x)[8], const Fft8::Tables &tb, int dir, cplx *b1, cplx *b2) {
    static cplx z[128][8], send[128][4];
    for (int t = 0; t < 128; t++) Fft8::pass_a(t, x[t], tb.ta, tb.twist[t], dir == 0, b1);
    for (int t = 0; t < 128; t++) Fft8::pass_b_first(t, b1, z[t], send[t]);
    for (int t = 0; t < 128; t++) Fft8::pass_b_second(t, z[t], send[t ^ 8], tb.tb[dir][t], b2);     // the shuffle: partner = lane ^ 8
    for (int t = 0; t < 128; t++) Fft8::pass_c(t, x[t], b2);
}
// forward transform of 1024 folded complex inputs (re, im interleaved); out[f] = sum_j x_j e^{i pi j/N} e^{-2 pi i jf/M}
extern "C" int emul_fft8_forward(const double *in, double *out) {
    static Fft8::Tables tb; fill_fft8_tables(tb);
    std::vector<cplx> b1(Fft8::BUF1), b2(Fft8::BUF2);
    static cplx x[128][8];
    for (int t = 0; t < 128; t++) {
        for (int p = 0; p < 8; p++) { int j = t + 128 * brev(p, 3); x[t][p] = cplx{in[2 * j], in[2 * j + 1]}; }
        Fft8::pre_twist(x[t]);
    }
    fft8_team(x, tb, 0, b1.data(), b2.data());
    for (int t = 0; t < 128; t++)
        for (int f2 = 0; f2 < 8; f2++) { out[2 * (128 * f2 + t)] = x[t][f2].re; out[2 * (128 * f2 + t) + 1] = x[t][f2].im; }
    return 0;
}
// inverse: X[1024] (interleaved) -> 2048 real coefficients (c[j] = re, c[j+M] = im of untwisted inverse DFT / M)
extern "C" int emul_fft8_inverse(const double *in, double *coef) {
    static Fft8::Tables tb; fill_fft8_tables(tb);
    std::vector<cplx> b1(Fft8::BUF1), b2(Fft8::BUF2);
    static cplx x[128][8];
    for (int t = 0; t < 128; t++)
        for (int p = 0; p < 8; p++) { int f = t + 128 * brev(p, 3); x[t][p] = cplx{in[2 * f], -in[2 * f + 1]}; }
    fft8_team(x, tb, 1, b1.data(), b2.data());
    for (int t = 0; t < 128; t++) {
        Fft8::post_twist(x[t]);
        for (int q = 0; q < 8; q++) { coef[t + 128 * q] = x[t][q].re; coef[t + 128 * q + 1024] = -x[t][q].im; }
    }
    return 0;
}

// full blind rotation with the latency kernel's stage sequence; the bootstrap key is converted by the standard routine
// (convert_bsk<.., 32, ..>), which pins that both FFTs agree on the Fourier-domain order
extern "C" int emul_blind_rotate_lat(int n, const uint64_t *bsk_std, const uint64_t *lwe, int batch, const uint64_t *lut,
                                     uint64_t *out_glwe) {
    typedef uint64_t T;
    typedef BRCore<T, 32, 1, 1, 23> C;
    constexpr int N = 2048, M = 1024, K1 = 2;
    static Fft8::Tables tb; fill_fft8_tables(tb);
    std::vector<cplx> fbsk;
    convert_bsk<T, 32, 1, 1, 23>(bsk_std, n, fbsk);
    std::vector<cplx> b1(Fft8::BUF1), b2(Fft8::BUF2), D(K1 * M);
    static cplx x[K1][128][8];
    for (int b = 0; b < batch; b++) {
        const uint64_t *ct = lwe + (size_t)b * (n + 1);
        std::vector<T> acc((size_t)K1 * N);
        int bhat = (int)(modswitch<T>(ct[n], C::LOGN) & (2 * N - 1));
        int a0 = (2 * N - bhat) & (2 * N - 1);
        for (int r = 0; r < K1; r++)
            for (int c = 0; c < N; c++) acc[(size_t)r * N + c] = rot_read<T, N>(&lut[(size_t)r * N], c, a0);
        for (int i = 0; i < n; i++) {
            int a = (int)(modswitch<T>(ct[i], C::LOGN) & (2 * N - 1));
            const cplx *ggsw = &fbsk[(size_t)i * C::GGSW_CPLX];
            for (int r = 0; r < K1; r++) {
                const T *acc_r = &acc[(size_t)r * N];
                for (int t = 0; t < 128; t++) {
                    for (int p = 0; p < 8; p++) {
                        int c = t + 128 * brev(p, 3);
                        T d0 = rot_read<T, N>(acc_r, c, a) - acc_r[c], d1 = rot_read<T, N>(acc_r, c + M, a) - acc_r[c + M];
                        x[r][t][p] = cplx{(double)C::single_level_digit(d0), (double)C::single_level_digit(d1)};
                    }
                    Fft8::pre_twist(x[r][t]);
                }
                fft8_team(x[r], tb, 0, b1.data(), b2.data());
                for (int t = 0; t < 128; t++)
                    for (int f2 = 0; f2 < 8; f2++) D[r * M + 128 * f2 + t] = x[r][t][f2];
            }
            for (int j = 0; j < K1; j++) {
                for (int t = 0; t < 128; t++)
                    for (int p = 0; p < 8; p++) {
                        int f = t + 128 * brev(p, 3);
                        cplx F = cmul(ggsw[(size_t)(j * K1 + j) * M + f], D[j * M + f]);            // own row first
                        C::mac_one(F, ggsw[(size_t)((1 - j) * K1 + j) * M + f], D[(1 - j) * M + f]);
                        x[j][t][p] = cplx{F.re, -F.im};
                    }
                fft8_team(x[j], tb, 1, b1.data(), b2.data());
                for (int t = 0; t < 128; t++) {
                    Fft8::post_twist(x[j][t]);
                    for (int q = 0; q < 8; q++) {
                        acc[(size_t)j * N + t + 128 * q] += torus_from_double<T>(x[j][t][q].re);
                        acc[(size_t)j * N + t + 128 * q + M] += torus_from_double<T>(-x[j][t][q].im);
                    }
                }
            }
        }
        for (size_t c = 0; c < acc.size(); c++) out_glwe[(size_t)b * K1 * N + c] = acc[c];
    }
    return 0;
}

// ---- multi-bit blind rotation (blind_rotate_mb.cu): same stage sequence, lanes executed one after another ------------
#include "br_multibit.cuh"
template <typename T, int R, int K, int L, int BL, int G>
static int run_mb(const uint64_t *bsk_std, int n, const uint64_t *lwe, int batch, const uint64_t *lut, uint64_t *out) {
    typedef MBCore<T, R, K, L, BL, G> C;
    static TwTables<R> tw; fill_tw_tables(tw);
    const int N = C::N, K1 = C::K1, NG = C::NG, ngroups = n / G;
    std::vector<cplx> fbsk;
    convert_bsk<T, R, K, L, BL>(bsk_std, ngroups * NG, fbsk);   // the key conversion is per GGSW polynomial
    cplx wr[R];
    for (int m = 0; m < R; m++) wr[m] = C::wr_entry(m);
    std::vector<T> acc((size_t)K1 * N), lutT((size_t)K1 * N), mask(n);
    std::vector<cplx> fbuf((size_t)K1 * L * C::REGION);
    static cplx x[R][R];
    static cplx F[K + 1][R][R];
    for (size_t c = 0; c < lutT.size(); c++) lutT[c] = (T)lut[c];
    for (int b = 0; b < batch; b++) {
        const uint64_t *ct = lwe + (size_t)b * (n + 1);
        for (int i = 0; i < n; i++) mask[i] = (T)ct[i];
        int bhat = (int)(modswitch<T>((T)ct[n], C::LOGN) & (2 * N - 1));
        int a0 = (2 * N - bhat) & (2 * N - 1);
        for (int r = 0; r < K1; r++)
            for (int c = 0; c < N; c++) acc[(size_t)r * N + c] = rot_read<T, N>(&lutT[(size_t)r * N], c, a0);
        for (int gi = 0; gi < ngroups; gi++) {
            uint16_t deg[C::NMONO];
            for (int idx = 1; idx < NG; idx++) deg[idx - 1] = C::group_degree(&mask[(size_t)gi * G], idx);
            for (int r = 0; r < K1; r++)
                for (int lane = 0; lane < R; lane++)
                    C::digits_plain_fwd_pass1(lane, &acc[(size_t)r * N], &fbuf[(size_t)r * L * C::REGION], tw.fwd);
            for (int r = 0; r < K1; r++)
                for (int t = 0; t < L; t++) {
                    cplx *reg = &fbuf[(size_t)(r * L + t) * C::REGION];
                    for (int lane = 0; lane < R; lane++) C::fwd_pass2_load(lane, x[lane], reg);
                    for (int lane = 0; lane < R; lane++) C::fwd_pass2_store(lane, x[lane], reg);
                }
            const cplx *group = &fbsk[(size_t)gi * C::GROUP_CPLX];
            for (int j = 0; j < K1; j++)
                for (int lane = 0; lane < R; lane++) C::mb_mac(lane, j, group, fbuf.data(), deg, wr, F[j][lane]);
            for (int j = 0; j < K1; j++)
                for (int lane = 0; lane < R; lane++) C::inv_pass1(lane, F[j][lane], &fbuf[(size_t)j * L * C::REGION], tw.inv);
            for (int j = 0; j < K1; j++)
                for (int lane = 0; lane < R; lane++) C::inv_pass2_store(lane, &fbuf[(size_t)j * L * C::REGION], &acc[(size_t)j * N]);
        }
        for (size_t c = 0; c < acc.size(); c++) out[(size_t)b * K1 * N + c] = (uint64_t)acc[c];
    }
    return 0;
}

extern "C" int emul_multibit_blind_rotate(int n, int k, int N, int base_log, int level, int bits, int grouping,
                                          const uint64_t *bsk_std, const uint64_t *lwe, int batch, const uint64_t *lut,
                                          uint64_t *out_glwe) {
#define CASE(T, R, K, L, BL, G) \
    if (bits == (int)sizeof(T) * 8 && N == 2 * R * R && k == K && level == L && base_log == BL && grouping == G) \
        return run_mb<T, R, K, L, BL, G>(bsk_std, n, lwe, batch, lut, out_glwe);
    CASE(uint64_t, 16, 1, 2, 12, 2)
    CASE(uint64_t, 16, 1, 2, 12, 3)
    CASE(uint64_t, 32, 1, 1, 22, 2)
    CASE(uint64_t, 32, 1, 1, 21, 3)
#undef CASE
    return 1;
}

// ---- device Fourier layout of one key polynomial (what bsk_convert writes): used to pin "position index == frequency index"
// against numpy's FFT, which is what lets a Fourier-domain key in natural order be uploaded without conversion --------------
extern "C" int emul_key_forward(int N, int bits, const uint64_t *poly, double *out) {
    if (N == 8192) {
        typedef BigCore<uint64_t, 1, 2, 15> C;
        constexpr int TH = 256, P = C::P;
        static BigTables tb; fill_big_tables(tb);
        std::vector<cplx> S1(C::M);
        static cplx x[TH][P];
        for (int t = 0; t < TH; t++) C::key_input(t, poly, x[t]);
        for (int t = 0; t < TH; t++) C::fwdA(t, x[t], S1.data(), tb.ta, tb.tq);
        for (int t = 0; t < TH; t++) C::fwdB_load(t, x[t], S1.data());
        for (int t = 0; t < TH; t++) C::fwdB_pass1(t, x[t], S1.data(), &tb);
        for (int t = 0; t < TH; t++) C::fwdB_pass2(t, x[t], S1.data());
        for (int t = 0; t < TH; t++) for (int p = 0; p < P; p++) { out[2 * (p * TH + t)] = x[t][p].re; out[2 * (p * TH + t) + 1] = x[t][p].im; }
        return 0;
    }
    std::vector<cplx> f;
    if (N == 2048 && bits == 64) convert_bsk<uint64_t, 32, 1, 1, 23>(poly, 1, f);
    else if (N == 512 && bits == 64) convert_bsk<uint64_t, 16, 3, 1, 18>(poly, 1, f);
    else if (N == 512 && bits == 32) convert_bsk<uint32_t, 16, 2, 3, 6>(poly, 1, f);
    else return 1;
    for (int i = 0; i < N / 2; i++) { out[2 * i] = f[i].re; out[2 * i + 1] = f[i].im; }   // first polynomial of the "key"
    return 0;
}

// twiddles.hpp -- host-side generation of the per-lane twiddle tables (long double, reduced angles).
#pragma once
#include "fft_reg.cuh"
#include "br_big.cuh"
#include "fft8.cuh"
#include <cmath>

namespace b200tfhe {

// e^{i pi num / N} with num reduced mod 2N before the long-double evaluation
inline cplx root_2N(long long num, int N) {
    long long m = ((num % (2LL * N)) + 2LL * N) % (2LL * N);
    const long double PI = 3.14159265358979323846264338327950288L;
    long double a = PI * (long double)m / (long double)N;
    return cplx{(double)cosl(a), (double)sinl(a)};
}

template <int R> void fill_tw_tables(TwTables<R> &t) {
    const int M = R * R, N = 2 * M;
    for (int lane = 0; lane < R; lane++) {
        for (int b = 0; b < 4; b++) {
            t.fwd[b][lane] = root_2N((long long)lane * (1 - 4 * b), N);          // u_l w_l^b
            cplx z = root_2N((long long)b * (4 * lane - 1), N);                  // z_l^b
            t.inv[b][lane] = cplx{z.re / M, z.im / M};                           // exact: M is a power of two
        }
        for (int a = 0; a < R / 4; a++) {
            t.fwd[4 + a][lane] = root_2N(-(long long)lane * 16 * a, N);          // w_l^{4a} = e^{-2 pi i l 4a/M}
            t.inv[4 + a][lane] = root_2N((long long)4 * a * (4 * lane - 1), N);  // z_l^{4a}
        }
    }
}

inline void fill_big_tables(BigTables &t) {
    const int N = 8192, M = 4096;
    (void)M;
    for (int b = 0; b < 4; b++)
        for (int th = 0; th < 256; th++) {
            t.ta[b][th] = root_2N((long long)th * (1 - 4 * b), N);   // u_t * w_4096^{t b}
            t.tq[b][th] = root_2N(-16LL * th * b, N);                // w_4096^{4 t b}
        }
    for (int lane = 0; lane < 16; lane++) {
        for (int b = 0; b < 4; b++) {
            cplx z = root_2N(-(long long)lane * b * 64, N);          // w_256^{lane b}
            t.pf[b][lane] = z; t.pi[b][lane] = cplx{z.re, -z.im};
            cplx q = root_2N(-(long long)lane * b * 256, N);         // w_256^{lane 4b}
            t.pf[4 + b][lane] = q; t.pi[4 + b][lane] = cplx{q.re, -q.im};
        }
    }
}

// tables of the 128-thread three-pass FFT (fft8.cuh)
inline void fill_fft8_tables(Fft8::Tables &t) {
    const int N = Fft8::N, M = Fft8::M;
    for (int tau = 0; tau < 128; tau++) t.twist[tau] = root_2N(tau, N);
    for (int g = 0; g < 8; g++)
        for (int u = 0; u < 16; u++) t.ta[g][u] = root_2N(-32LL * u * g, N);
    for (int beta = 0; beta < 128; beta++) {
        const long long tau0 = beta & 7, half = (beta >> 3) & 1, g = beta >> 4, c0 = g + 32 * half;
        Fft8::TwB &f = t.tb[0][beta], &v = t.tb[1][beta];
        f.a = root_2N(-4 * tau0 * c0, N);
        f.b1 = root_2N(-4 * tau0 * 8, N); f.b2 = root_2N(-4 * tau0 * 16, N); f.b3 = root_2N(-4 * tau0 * 24, N);
        f.c = root_2N(-4 * tau0 * 64, N);
        cplx z = root_2N(-4 * tau0 * c0 + c0, N);
        v.a = cplx{z.re / M, z.im / M};    // exact: M is a power of two
        v.b1 = root_2N(-4 * tau0 * 8 + 8, N); v.b2 = root_2N(-4 * tau0 * 16 + 16, N); v.b3 = root_2N(-4 * tau0 * 24 + 24, N);
        v.c = root_2N(-4 * tau0 * 64 + 64, N);
    }
}

}  // namespace b200tfhe

// twiddles.hpp -- host-side generation of the per-lane twiddle tables (long double, reduced angles).
#pragma once
#include "fft_reg.cuh"
#include "br_big.cuh"
#include "fft3.cuh"
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

// tables of the 64-thread three-pass FFT (fft3.cuh)
inline void fill_fft3_tables(Fft3::Tables &t) {
    const int N = Fft3::N, M = Fft3::M;
    for (int g2 = 0; g2 < 16; g2++) {
        for (int tau = 0; tau < 64; tau++) t.ta_fwd[g2][tau] = root_2N((long long)tau - 64LL * (tau >> 4) * g2, N);
        for (int u = 0; u < 4; u++) t.ta_inv[g2][u] = root_2N(-64LL * u * g2, N);
    }
    for (int beta = 0; beta < 64; beta++) {
        const int j1 = beta & 15, gq = beta >> 4;
        for (int e = 0; e < 4; e++)
            for (int g1 = 0; g1 < 4; g1++) {
                const long long f2 = 16 * g1 + 4 * gq + e;
                t.tb[0][e * 4 + g1][beta] = root_2N(-4LL * j1 * f2, N);
                cplx z = root_2N(-4LL * j1 * f2 + f2, N);
                t.tb[1][e * 4 + g1][beta] = cplx{z.re / M, z.im / M};    // exact: M is a power of two
            }
    }
}

}  // namespace b200tfhe

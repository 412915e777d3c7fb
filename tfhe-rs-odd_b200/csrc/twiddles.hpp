// twiddles.hpp -- host-side generation of the per-lane twiddle tables (long double, reduced angles).
#pragma once
#include "fft_reg.cuh"
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

}  // namespace b200tfhe

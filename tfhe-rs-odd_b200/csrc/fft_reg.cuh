// fft_reg.cuh -- register-resident radix-R complex FFT and the two-pass "team" FFT of size M = R*R.
//
// Replaces, for the B200 build, the reference's FftView::{forward_as_integer, forward_as_torus,
// add_backward_in_place_as_torus} (tfhe/src/core_crypto/fft_impl/fft64/math/fft/mod.rs:197-326,496-557)
// and the un-vendored concrete-fft 0.3.0 `unordered::Plan::{fwd,inv}` they call.  Same mathematics
// (fold N reals to M = N/2 complex, twist by e^{i pi j/N}, complex DFT of size M, kernel e^{-2 pi i jk/M}),
// different algorithm: a team of R threads holds R points each, does a radix-R DIT entirely in
// registers with compile-time twiddles, transposes once through shared memory, and does a second
// radix-R DIT.  The Fourier-domain coefficient order is the device's own ("position" p*R + lane holds
// frequency lane + R*p); the bootstrap key is converted with the same routine so pointwise products agree,
// exactly as concrete-fft's `unordered` plans do on the CPU.
//
// Everything is __host__ __device__ so tests/native/emul.cpp can execute the same code lane by lane.
#pragma once
#include "common.cuh"
#include <utility>

namespace b200tfhe {

struct alignas(16) cplx { double re, im; };

// cos(2 pi q / 128), q = 0..32, correctly rounded (generated with mpmath at 200 bits, tools/gen_tables.py)
constexpr double COS128_TAB[33] = {
    1.0, 0.9987954562051724, 0.9951847266721969, 0.989176509964781, 0.9807852804032304, 0.970031253194544,
    0.9569403357322088, 0.9415440651830208, 0.9238795325112867, 0.9039892931234433, 0.881921264348355,
    0.8577286100002721, 0.8314696123025452, 0.8032075314806449, 0.773010453362737, 0.7409511253549591,
    0.7071067811865476, 0.6715589548470184, 0.6343932841636455, 0.5956993044924334, 0.5555702330196022,
    0.5141027441932218, 0.47139673682599764, 0.4275550934302821, 0.3826834323650898, 0.33688985339222005,
    0.2902846772544624, 0.2429801799032639, 0.19509032201612828, 0.14673047445536175, 0.0980171403295606,
    0.049067674327418015, 0.0};

HDC int mod128(int q) { return ((q % 128) + 128) % 128; }
HDC double cos128(int q) {
    q = mod128(q);
    if (q > 64) q = 128 - q;
    return q > 32 ? -COS128_TAB[64 - q] : COS128_TAB[q];
}
HDC double sin128(int q) { return cos128(q - 32); }

HDC int brev(int v, int bits) {
    int r = 0;
    for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
    return r;
}
HDC int clog2(int v) { int l = 0; while ((1 << l) < v) l++; return l; }

template <int I, int E, class F> HD void static_for(F &&f) {
    if constexpr (I < E) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, E>(f);
    }
}

HD double fma_(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}

// x * e^{2 pi i Q/128}, Q compile-time
template <int Q> HD cplx cmul_root(cplx x) {
    constexpr int q = mod128(Q);
    if constexpr (q == 0) return x;
    else if constexpr (q == 32) return cplx{-x.im, x.re};
    else if constexpr (q == 64) return cplx{-x.re, -x.im};
    else if constexpr (q == 96) return cplx{x.im, -x.re};
    else {
        constexpr double c = cos128(q), s = sin128(q);
        return cplx{fma_(-s, x.im, c * x.re), fma_(s, x.re, c * x.im)};
    }
}

// 16-byte load of read-only key data through the non-coherent path.  It allocates in L1 on purpose: the ciphertexts of
// a CTA read the same key elements within a few hundred cycles of each other, so all but the first hit L1 (measured on
// the TMEM kernel: 76.5k vs 70.9k PBS/s with L1::no_allocate).
HD cplx ldg_stream(const cplx *p) {
#if defined(__CUDA_ARCH__)
    cplx v;
    asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(v.re), "=d"(v.im) : "l"(p));
    return v;
#else
    return *p;
#endif
}

#if defined(__CUDA_ARCH__)
DEV void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
#endif

HD cplx cmul(cplx a, cplx b) { return cplx{fma_(-a.im, b.im, a.re * b.re), fma_(a.im, b.re, a.re * b.im)}; }

// DIT butterfly (a, b) <- (a + W b, a - W b), W = e^{2 pi i Q/128}.  General twiddles use the
// factored form W b = c (b + i t b) (t = tan) so the butterfly is 6 FMAs instead of 8 flops-instructions.
template <int Q> HD void bfly(cplx &a, cplx &b) {
    constexpr int q = mod128(Q);
    if constexpr (q == 0 || q == 32 || q == 64 || q == 96) {
        cplx t = cmul_root<q>(b);
        b = cplx{a.re - t.re, a.im - t.im};
        a = cplx{a.re + t.re, a.im + t.im};
    } else {
        constexpr double c = cos128(q), s = sin128(q);
        constexpr bool use_tan = (c < 0 ? -c : c) >= (s < 0 ? -s : s);
        if constexpr (use_tan) {
            constexpr double tn = s / c;
            double t1 = fma_(-tn, b.im, b.re), t2 = fma_(tn, b.re, b.im);
            b = cplx{fma_(-c, t1, a.re), fma_(-c, t2, a.im)};
            a = cplx{fma_(c, t1, a.re), fma_(c, t2, a.im)};
        } else {
            constexpr double ct = c / s;
            double t1 = fma_(ct, b.re, -b.im), t2 = fma_(ct, b.im, b.re);
            b = cplx{fma_(-s, t1, a.re), fma_(-s, t2, a.im)};
            a = cplx{fma_(s, t1, a.re), fma_(s, t2, a.im)};
        }
    }
}

// In-register radix-R decimation-in-time FFT.  Input: x[p] holds element brev(p); output natural order.
// DIR = -1: kernel e^{-2 pi i jk/R} (forward); DIR = +1: unnormalised inverse.
template <int R, int DIR> HD void fft_dit(cplx (&x)[R]) {
    static_for<1, clog2(R) + 1>([&](auto s) {
        constexpr int len = 1 << decltype(s)::value, half = len / 2;
        static_for<0, R / len>([&](auto bi) {
            static_for<0, half>([&](auto q) {
                constexpr int base = decltype(bi)::value * len + decltype(q)::value;
                bfly<DIR * decltype(q)::value * (128 / len)>(x[base], x[base + half]);
            });
        });
    });
}

// Per-lane twiddle tables of the team FFT (filled on the host in long double, twiddles.hpp):
//   fwd_r[b][lane] = u_lane * w_lane^b      (b = 0..3)      w_l = e^{-2 pi i l/M}, u_l = e^{i pi l/N}
//   fwd_q[a][lane] = w_lane^{4a}            (a = 0..R/4-1)
//   inv_r[b][lane] = z_lane^b / M           (b = 0..3)      z_l = e^{+2 pi i l/M} e^{-i pi/N}
//   inv_q[a][lane] = z_lane^{4a}            (a = 0..R/4-1)
// so that the middle twiddle of register index e is r[e&3] * q[e>>2] (one multiply of two correctly
// rounded values).  The twist factor that depends on the register index is a compile-time root.
template <int R> struct TwTables {
    static constexpr int ENTRIES = 4 + R / 4;
    cplx fwd[ENTRIES][R];
    cplx inv[ENTRIES][R];
};

template <int R> struct TeamFFT {
    static constexpr int M = R * R, N = 2 * M, LOGR = clog2(R);
    static constexpr int ROW = R + 1;            // padded row stride (cplx units): conflict-free transposes
    static constexpr int REGION = R * ROW;       // cplx per exchange / Fourier buffer
    static constexpr int TWQ = 32 / R;           // twist step in 128ths of a turn: e^{i pi m/(2R)} = e^{2 pi i m TWQ/128}
    static_assert(R == 16 || R == 32, "team FFT sizes: M = 256 or 1024");

    // ---- forward ----------------------------------------------------------------------------------
    // x[p] must hold the folded input point m = brev(p) of this lane (coefficient lane + R*m as real part,
    // + M as imaginary part), NOT yet twisted.  After the call x[k1] = u_l w_l^{k1} sum_m x_m v_m w_R^{m k1},
    // and it has been written to S[k1][lane].
    HD static void fwd_pass1_compute(int lane, cplx (&x)[R], const cplx (*tw)[R]) {
        twist(x);
        fwd_pass1_compute_twisted(lane, x, tw);
    }
    // the two halves of fwd_pass1_compute, for kernels that twist in another warp (blind_rotate_ws.cu)
    HD static void twist(cplx (&x)[R]) {
        static_for<0, R>([&](auto p) {
            constexpr int m = brev(decltype(p)::value, LOGR);
            x[decltype(p)::value] = cmul_root<m * TWQ>(x[decltype(p)::value]);
        });
    }
    HD static void fwd_pass1_compute_twisted(int lane, cplx (&x)[R], const cplx (*tw)[R]) {
        fft_dit<R, -1>(x);
        mid_twiddle(lane, x, tw);
    }
    HD static void pass1_store(int lane, const cplx (&x)[R], cplx *region) {
        static_for<0, R>([&](auto k1) { region[decltype(k1)::value * ROW + lane] = x[decltype(k1)::value]; });
    }
    HD static void fwd_pass1(int lane, cplx (&x)[R], cplx *region, const cplx (*tw)[R]) {
        fwd_pass1_compute(lane, x, tw);
        pass1_store(lane, x, region);
    }
    // x[e] *= r[e & 3] * q[e >> 2]  (per-lane tables, see TwTables)
    HD static void mid_twiddle(int lane, cplx (&x)[R], const cplx (*tw)[R]) {
        cplx r[4];
        for (int b = 0; b < 4; b++) r[b] = tw[b][lane];
        static_for<0, R / 4>([&](auto a) {
            if constexpr (decltype(a)::value == 0) {
                static_for<0, 4>([&](auto b) { x[decltype(b)::value] = cmul(x[decltype(b)::value], r[decltype(b)::value]); });
            } else {
                cplx q = tw[4 + decltype(a)::value][lane];
                static_for<0, 4>([&](auto b) {
                    constexpr int e = decltype(a)::value * 4 + decltype(b)::value;
                    x[e] = cmul(x[e], cmul(r[decltype(b)::value], q));
                });
            }
        });
    }
    // team sync, then: x[p] <- S[lane][brev(p)]
    HD static void fwd_pass2_load(int lane, cplx (&x)[R], const cplx *region) {
        static_for<0, R>([&](auto p) { x[decltype(p)::value] = region[lane * ROW + brev(decltype(p)::value, LOGR)]; });
    }
    // team sync, then DIT; x[k2] = X[lane + R*k2]; position of that value in a Fourier buffer = k2*R + lane
    HD static void fwd_pass2_compute(cplx (&x)[R]) { fft_dit<R, -1>(x); }

    // ---- inverse ----------------------------------------------------------------------------------
    // x[p] must hold Fourier position brev(p)*R + lane.  Writes S'[l][lane] = conj-twiddled partial result.
    HD static void inv_pass1_compute(int lane, cplx (&x)[R], const cplx (*tw)[R]) {
        fft_dit<R, +1>(x);
        mid_twiddle(lane, x, tw);
    }
    HD static void inv_pass1(int lane, cplx (&x)[R], cplx *region, const cplx (*tw)[R]) {
        inv_pass1_compute(lane, x, tw);
        pass1_store(lane, x, region);
    }
    // team sync, then: x[p] <- S'[lane][brev(p)], DIT, un-twist by the compile-time root.
    // Result: x[m] = (1/M) * conj(twist) * inverse DFT, i.e. re -> coefficient lane + R*m, im -> that + M.
    HD static void inv_pass2(int lane, cplx (&x)[R], const cplx *region) {
        static_for<0, R>([&](auto p) { x[decltype(p)::value] = region[lane * ROW + brev(decltype(p)::value, LOGR)]; });
        fft_dit<R, +1>(x);
        static_for<0, R>([&](auto m) { x[decltype(m)::value] = cmul_root<-decltype(m)::value * TWQ>(x[decltype(m)::value]); });
    }
};

// ---- torus <-> double -------------------------------------------------------------------------------
// Small signed integer (|d| < 2^31) to double without the conversion pipe: 2^52+2^31 magic.
HD double int_to_double(int32_t d) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double(0x43300000, (int)((uint32_t)d ^ 0x80000000u)) - 4503601774854144.0;
#else
    return (double)d;
#endif
}

// same, divided by 2^SHIFT (exact): lets a kernel fold the 1/M of the inverse transform into the digit conversion
template <int SHIFT> HD double int_to_double_scaled(int32_t d) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double(0x43300000 - (SHIFT << 20), (int)((uint32_t)d ^ 0x80000000u)) - 4503601774854144.0 / (double)(1 << SHIFT);
#else
    return (double)d / (double)(1 << SHIFT);
#endif
}

// from_torus (commons/math/torus/mod.rs:72-78) followed by the signed wrapping cast (fft/mod.rs:285-304)
template <typename T> HD T torus_from_double(double t) {
    const double MAGIC = 6755399441055744.0;  // 1.5 * 2^52 : round to nearest integer for |t| < 2^51
    double r = (t + MAGIC) - MAGIC;
    double f = t - r;
    if constexpr (TorusTraits<T>::BITS == 64) {
        f *= 18446744073709551616.0;
#if defined(__CUDA_ARCH__)
        return (uint64_t)__double2ll_rn(f);
#else
        if (f >= 9223372036854775807.0) return (uint64_t)INT64_MAX;
        if (f <= -9223372036854775808.0) return (uint64_t)INT64_MIN;
        return (uint64_t)(int64_t)__builtin_rint(f);
#endif
    } else {
        f *= 4294967296.0;
#if defined(__CUDA_ARCH__)
        return (uint32_t)__double2int_rn(f);
#else
        if (f >= 2147483647.0) return (uint32_t)INT32_MAX;
        return (uint32_t)(int32_t)__builtin_rint(f);
#endif
    }
}

// signed torus value -> double in [-0.5, 0.5): forward_as_torus normalisation (fft/mod.rs:197-218)
template <typename T> HD double torus_to_double(T v) {
    if constexpr (TorusTraits<T>::BITS == 64) return (double)(int64_t)v * 5.421010862427522e-20;  // 2^-64
    else return (double)(int32_t)v * 2.3283064365386963e-10;                                      // 2^-32
}

}  // namespace b200tfhe

// common.cuh -- parameter block and the bit-exact integer primitives shared by every kernel.
//
// Reference semantics restated (citations relative to /root/reference/tfhe/src/):
//   closest_representable / decompose   core_crypto/commons/math/decomposition/decomposer.rs:99-119, iter.rs:120-127
//   tensor decomposition                core_crypto/fft_impl/fft64/math/decomposition.rs:27-86
//   fast_pbs_modulus_switch             core_crypto/fft_impl/common.rs:26-43
// All functions are __host__ __device__ so that tests/native/emul.cpp can run the same code on the CPU.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define HD __host__ __device__ __forceinline__
#define DEV __device__ __forceinline__
#define HDC __host__ __device__ constexpr
#else
#define HDC constexpr
#define HD inline
#define DEV inline
#endif

namespace b200tfhe {

// Crypto parameters of one KS+PBS pipeline.
struct Params {
    int n;        // small LWE dimension
    int k;        // GLWE dimension
    int N;        // polynomial size
    int pbs_base_log, pbs_level;
    int ks_base_log, ks_level;
    int bits;     // 32 or 64: torus width
};

template <typename T> struct TorusTraits;
template <> struct TorusTraits<uint64_t> { static constexpr int BITS = 64; typedef int64_t Signed; };
template <> struct TorusTraits<uint32_t> { static constexpr int BITS = 32; typedef int32_t Signed; };

// ((x >> (W - log2N - 2)) + 1) >> 1  in [0, 2N]
template <typename T> HD uint32_t modswitch(T x, int log2N) {
    constexpr int W = TorusTraits<T>::BITS;
    T v = x >> (W - log2N - 2);
    v += 1;
    v >>= 1;
    return (uint32_t)v;
}

// state = closest_representable(x) >> (W - b*l), masked to b*l bits (the reference's `<< shift` wraps).
template <typename T> HD T decomp_init_state(T x, int base_log, int level) {
    constexpr int W = TorusTraits<T>::BITS;
    const int r = W - base_log * level;               // non-representable bits, >= 1 for every supported set
    T v = x >> (r - 1);
    v += 1;
    v >>= 1;
    const T m = (base_log * level == W) ? ~(T)0 : (((T)1 << (base_log * level)) - 1);
    return v & m;
}

// One step of the balanced digit recurrence; returns the digit as a signed value in [-B/2, B/2].
template <typename T> HD typename TorusTraits<T>::Signed decomp_next_digit(T &state, int base_log) {
    const T mod_b_mask = ((T)1 << base_log) - 1;
    T res = state & mod_b_mask;
    state >>= base_log;
    T carry = ((res - 1) | state) & res;
    carry >>= (base_log - 1);
    state += carry;
    return (typename TorusTraits<T>::Signed)(res - (carry << base_log));
}

// register type of a decomposition state of STATE_BITS = level * base_log bits: the recurrence never leaves
// [0, 2^STATE_BITS] (masked at init, + one carry), so 32 bits do whenever STATE_BITS <= 31
template <typename T, int STATE_BITS, bool NARROW = (STATE_BITS <= 31)> struct StateWord { typedef T type; };
template <typename T, int STATE_BITS> struct StateWord<T, STATE_BITS, true> { typedef uint32_t type; };

HD int ilog2(int x) { int l = 0; while ((1 << l) < x) l++; return l; }

}  // namespace b200tfhe

// wopbs_kernels.cu -- the integer steps between the keyswitch and the bootstrap of every round of the bit extraction
// (first half of the WoP-PBS).
//
// Reference: extract_bits, tfhe/src/core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs:66-225 (public wrapper
// extract_bits_from_lwe_ciphertext_mem_optimized, core_crypto/algorithms/lwe_wopbs.rs:321-364).  Round `bit_idx` (from the
// least significant extracted bit upwards) of one ciphertext is
//     shifted   = buffer << (W - delta_log - bit_idx - 1)                      :160-166
//     ks        = keyswitch(shifted);  out[nbits-1-bit_idx] = ks                :169-184
//     ks.body  += q/4;  pbs = bootstrap(ks, LUT = constant -2^(delta_log-1+bit_idx))   :193-213
//     pbs.body += 2^(delta_log+bit_idx-1);  buffer -= pbs                       :217-223
// The keyswitch and the bootstrap are the batched kernels of the hot path; the two kernels here do the wrapping-integer
// element-wise steps for the whole batch, fused so that a round is 2 small launches + KS + PBS and nothing returns to the host.
#include "kernels.h"

namespace b200tfhe {

// buffer -= pbs (+ alpha on the body) for rounds after the first, then shifted = buffer << sh
template <typename T>
__global__ void __launch_bounds__(256)
eb_update_shift_kernel(T *__restrict__ buffer, const T *__restrict__ pbs, T *__restrict__ shifted, size_t len, size_t total, int sh,
                       T alpha) {
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    T v = buffer[e];
    if (pbs) {
        T s = pbs[e];
        if (e % len == len - 1) s += alpha;
        v -= s;
        buffer[e] = v;
    }
    shifted[e] = (T)(v << sh);
}

// out[b][slot] = ks[b];  ks[b].body += q/4
template <typename T>
__global__ void __launch_bounds__(256)
eb_store_kernel(T *__restrict__ ks, T *__restrict__ out, size_t len, size_t total, int nbits, int slot, int add_quarter) {
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    const size_t b = e / len, c = e % len;
    const T v = ks[e];
    out[(b * nbits + slot) * len + c] = v;
    if (add_quarter && c == len - 1) ks[e] = v + ((T)1 << (sizeof(T) * 8 - 2));
}

int eb_update_shift_launch(void *buffer, const void *pbs, void *shifted, size_t len, int batch, int sh, uint64_t alpha, int bits,
                           cudaStream_t st) {
    const size_t total = (size_t)batch * len;
    if (total == 0) return 0;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64)
        eb_update_shift_kernel<uint64_t><<<grid, 256, 0, st>>>((uint64_t *)buffer, (const uint64_t *)pbs, (uint64_t *)shifted, len, total, sh, alpha);
    else
        eb_update_shift_kernel<uint32_t><<<grid, 256, 0, st>>>((uint32_t *)buffer, (const uint32_t *)pbs, (uint32_t *)shifted, len, total, sh,
                                                                 (uint32_t)alpha);
    return (int)cudaGetLastError();
}

int eb_store_launch(void *ks, void *out, size_t len, int batch, int nbits, int slot, int add_quarter, int bits, cudaStream_t st) {
    const size_t total = (size_t)batch * len;
    if (total == 0) return 0;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64) eb_store_kernel<uint64_t><<<grid, 256, 0, st>>>((uint64_t *)ks, (uint64_t *)out, len, total, nbits, slot, add_quarter);
    else eb_store_kernel<uint32_t><<<grid, 256, 0, st>>>((uint32_t *)ks, (uint32_t *)out, len, total, nbits, slot, add_quarter);
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

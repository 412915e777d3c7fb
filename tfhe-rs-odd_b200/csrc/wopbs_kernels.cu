// wopbs_kernels.cu -- the element-wise integer steps of the WoP-PBS: between the keyswitch and the bootstrap of
// every round of the bit extraction, and around the bootstraps of the circuit bootstrap (further down).
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

// ---- circuit bootstrap (wop_pbs/mod.rs:243-344, 369-444) -------------------------------------------------------------------
// Entry e = (ct, i, lvl) of the bootstrap batch: the input bit nbits-1-i of ciphertext ct (the vertical-packing blind rotation
// consumes the GGSW list from its LAST element, :874, so the list is built reversed), body + q/4 (:405-407; delta_log = W-1 makes
// the cleartext multiplication of :398-402 a multiplication by 1), accumulator index lvl.
template <typename T>
__global__ void __launch_bounds__(256)
cbs_prep_kernel(const T *__restrict__ bits_in, T *__restrict__ out, uint32_t *__restrict__ lut_idx, size_t len, size_t total, int nbits,
                int levels) {
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    const size_t entry = e / len, c = e % len;
    const int lvl = (int)(entry % levels);
    const size_t ci = entry / levels, ct = ci / nbits;
    const int i = (int)(ci % nbits);
    T v = bits_in[(ct * nbits + (nbits - 1 - i)) * len + c];
    if (c == len - 1) {
        v += (T)1 << (sizeof(T) * 8 - 2);
        lut_idx[entry] = (uint32_t)lvl;
    }
    out[e] = v;
}

// body of bootstrap output e += 2^(W - 1 - base_log * (lvl + 1)),  lvl = e % levels   (:441-443)
template <typename T>
__global__ void __launch_bounds__(256) cbs_post_kernel(T *__restrict__ pbs, size_t len, size_t entries, int levels, int base_log) {
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (e >= entries) return;
    const int lvl = (int)(e % levels);
    pbs[e * len + len - 1] += (T)1 << (sizeof(T) * 8 - 1 - base_log * (lvl + 1));
}

__global__ void __launch_bounds__(256) mod_index_kernel(uint32_t *__restrict__ idx, size_t total, uint32_t mod) {
    const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (e < total) idx[e] = (uint32_t)(e % mod);
}

int cbs_prep_launch(const void *bits_in, void *out, uint32_t *lut_idx, size_t len, int batch, int nbits, int levels, int bits, cudaStream_t st) {
    const size_t total = (size_t)batch * nbits * levels * len;
    if (total == 0) return 0;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64) cbs_prep_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)bits_in, (uint64_t *)out, lut_idx, len, total, nbits, levels);
    else cbs_prep_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)bits_in, (uint32_t *)out, lut_idx, len, total, nbits, levels);
    return (int)cudaGetLastError();
}

int cbs_post_launch(void *pbs, size_t len, size_t entries, int levels, int base_log, int bits, cudaStream_t st) {
    if (entries == 0) return 0;
    const unsigned grid = (unsigned)((entries + 255) / 256);
    if (bits == 64) cbs_post_kernel<uint64_t><<<grid, 256, 0, st>>>((uint64_t *)pbs, len, entries, levels, base_log);
    else cbs_post_kernel<uint32_t><<<grid, 256, 0, st>>>((uint32_t *)pbs, len, entries, levels, base_log);
    return (int)cudaGetLastError();
}

int mod_index_launch(uint32_t *idx, size_t total, uint32_t mod, cudaStream_t st) {
    if (total == 0) return 0;
    mod_index_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(idx, total, mod);
    return (int)cudaGetLastError();
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

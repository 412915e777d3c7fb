// wave_kernels.cu -- device-resident "waves" of shortint / boolean operations over a ciphertext pool.
//
// A wave is a list of independent operations  pool[dst] <- LUT( sum_k scale_k * pool[src_k] + plaintext )  (LUT optional):
// exactly the shape of the reference's leveled prologues followed by a bootstrap --
//   * bivariate PBS: unchecked_scalar_mul_assign + unchecked_add_assign, then apply_lookup_table
//     (shortint/server_key/bivariate_pbs.rs:167-182),
//   * unchecked_add / scalar ops (core_crypto/algorithms/lwe_linear_algebra.rs:191-215, 276-290, 556-572),
//   * message_extract / carry_extract (shortint/server_key/mod.rs:539-545, 619-623).
// Every operation of a wave reads the pool as it was BEFORE the wave (gather into a staging list, bootstrap the staged
// list with the batched KS+PBS kernels, scatter), so a dependent computation (integer carry propagation, Trivium rounds)
// runs as a sequence of waves without a ciphertext ever leaving HBM.
#include "kernels.h"

namespace b200tfhe {

template <typename T>
__global__ void wave_gather_kernel(const T *__restrict__ pool, size_t len, const WaveOp *__restrict__ ops, int n_ops,
                                   T *__restrict__ staging, uint32_t *__restrict__ lut_idx, int n_pbs) {
    const int op = blockIdx.x;      // ops on grid.x (up to 2^31 - 1 per wave; grid.y is capped at 65535), tiles of a slot on grid.y
    if (op >= n_ops) return;
    const WaveOp o = ops[op];
    if (op < n_pbs && blockIdx.y == 0 && threadIdx.x == 0) lut_idx[op] = (uint32_t)o.lut;
    const T *s0 = pool + (size_t)o.src[0] * len, *s1 = pool + (size_t)o.src[1] * len;
    const T *s2 = pool + (size_t)o.src[2] * len, *s3 = pool + (size_t)o.src[3] * len, *s4 = pool + (size_t)o.src4 * len;
    const T c0 = (T)o.scale[0], c1 = (T)o.scale[1], c2 = (T)o.scale[2], c3 = (T)o.scale[3], c4 = (T)(int64_t)o.scale4;
    T *dst = staging + (size_t)op * len;
    for (size_t e = (size_t)blockIdx.y * blockDim.x + threadIdx.x; e < len; e += (size_t)gridDim.y * blockDim.x) {
        T v = 0;                              // wrapping arithmetic on the torus; a zero scale skips the term
        if (c0) v += c0 * s0[e];
        if (c1) v += c1 * s1[e];
        if (c2) v += c2 * s2[e];
        if (c3) v += c3 * s3[e];
        if (c4) v += c4 * s4[e];
        if (e == len - 1) v += (T)o.plaintext;
        dst[e] = v;
    }
}

template <typename T>
__global__ void wave_scatter_kernel(T *__restrict__ pool, size_t len, const WaveOp *__restrict__ ops, int n_ops, int n_pbs,
                                    const T *__restrict__ pbs_out, const T *__restrict__ staging) {
    const int op = blockIdx.x;
    if (op >= n_ops) return;
    const T *src = (op < n_pbs ? pbs_out : staging) + (size_t)op * len;
    T *dst = pool + (size_t)ops[op].dst * len;
    for (size_t e = (size_t)blockIdx.y * blockDim.x + threadIdx.x; e < len; e += (size_t)gridDim.y * blockDim.x) dst[e] = src[e];
}

int wave_gather_launch(const void *pool, size_t len, const WaveOp *d_ops, int n_ops, void *staging, uint32_t *d_lut_idx, int n_pbs,
                       int bits, cudaStream_t st) {
    if (n_ops <= 0) return 0;
    dim3 grid((unsigned)n_ops, (unsigned)((len + 255) / 256 > 8 ? 8 : (len + 255) / 256));
    if (bits == 64) wave_gather_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)pool, len, d_ops, n_ops, (uint64_t *)staging, d_lut_idx, n_pbs);
    else wave_gather_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)pool, len, d_ops, n_ops, (uint32_t *)staging, d_lut_idx, n_pbs);
    return (int)cudaGetLastError();
}

int wave_scatter_launch(void *pool, size_t len, const WaveOp *d_ops, int n_ops, int n_pbs, const void *pbs_out, const void *staging,
                        int bits, cudaStream_t st) {
    if (n_ops <= 0) return 0;
    dim3 grid((unsigned)n_ops, (unsigned)((len + 255) / 256 > 8 ? 8 : (len + 255) / 256));
    if (bits == 64) wave_scatter_kernel<uint64_t><<<grid, 256, 0, st>>>((uint64_t *)pool, len, d_ops, n_ops, n_pbs, (const uint64_t *)pbs_out, (const uint64_t *)staging);
    else wave_scatter_kernel<uint32_t><<<grid, 256, 0, st>>>((uint32_t *)pool, len, d_ops, n_ops, n_pbs, (const uint32_t *)pbs_out, (const uint32_t *)staging);
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

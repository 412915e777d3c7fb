// keyswitch.cu -- batched LWE keyswitch, bit-exact wrapping integer arithmetic.
//
// Replaces keyswitch_lwe_ciphertext (tfhe/src/core_crypto/algorithms/lwe_keyswitch.rs:96-170; inner op
// slice_wrapping_sub_scalar_mul_assign, slice_algorithms.rs:363-459):
//     out = (0,...,0,b_in) - sum_i sum_t digit_t(a_i) * KSK[i][t]
// with the key kept in the reference's own layout [n_in][l_ks][n_out+1] (levels stored l_ks -> 1, the order
// the decomposition iterator emits them, lwe_keyswitch_key_generation.rs:109-120).
//
// Batched, this is an integer "GEMM" over Z/2^W:  [batch x (n_in*l)] digits  x  [(n_in*l) x (n_out+1)] key.
// Tiling: a CTA owns BT ciphertexts x JT output columns and walks all key rows; a thread owns a 4x4
// register tile.  Digits are made non-negative (d' = d + B/2 in [0, B]) so one multiply-add is
// IMAD.WIDE.U32 + IMAD (u64) or a single IMAD (u32); the offset is undone with a per-key column sum:
//     sum (d' - B/2) K = sum d' K - (B/2) * colsum,   colsum[j] = sum_{i,t} KSK[i][t][j]   (computed at upload).
// All arithmetic wraps mod 2^W, so the result is bit-identical to the serial reference loop.
#include "kernels.h"

namespace b200tfhe {

constexpr int KS_BT = 32;     // ciphertexts per CTA
constexpr int KS_JT = 128;    // output columns per CTA
constexpr int KS_IC = 32;     // input coefficients per digit chunk
constexpr int KS_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(KS_THREADS)
ks_kernel(const T *__restrict__ in, size_t in_stride, T *__restrict__ out, size_t out_stride,
          const T *__restrict__ ksk, const T *__restrict__ colsum, int n_in, int n_out, int level, int base_log,
          int batch, int body_col) {
    // digits of the current chunk: dig[(ii*level + t)][b] as bytes, 4 ciphertexts packed per 32-bit word
    extern __shared__ __align__(16) unsigned char dig[];
    const int tid = threadIdx.x;
    const int cg = tid % 32, bg = tid / 32;          // column group (4 columns), ciphertext group (4 cts)
    const int b0 = blockIdx.x * KS_BT, j0 = blockIdx.y * KS_JT + cg * 4;
    const int width = n_out + 1;
    const uint32_t half_b = 1u << (base_log - 1);

    T acc[4][4];
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
        for (int y = 0; y < 4; y++) acc[x][y] = 0;

    const bool col_ok[4] = {j0 + 0 < width, j0 + 1 < width, j0 + 2 < width, j0 + 3 < width};
    const bool vec_ok = (j0 + 3 < width) && ((((size_t)width * sizeof(T)) % 16) == 0);

    for (int i0 = 0; i0 < n_in; i0 += KS_IC) {
        __syncthreads();
        // ---- decompose KS_BT x KS_IC input coefficients into `level` digits each ----
        for (int e = tid; e < KS_BT * KS_IC; e += KS_THREADS) {
            const int b = e / KS_IC, ii = e % KS_IC;
            const int gb = b0 + b, gi = i0 + ii;
            T v = (gb < batch && gi < n_in) ? in[(size_t)gb * in_stride + gi] : (T)0;
            T state = decomp_init_state<T>(v, base_log, level);
            for (int t = 0; t < level; t++) {
                int d = (int)decomp_next_digit<T>(state, base_log);
                dig[(size_t)(ii * level + t) * KS_BT + b] = (unsigned char)(d + (int)half_b);
            }
        }
        __syncthreads();
        // ---- rows of this chunk ----
        const int rows = min(KS_IC, n_in - i0) * level;
        const T *krow = ksk + (size_t)i0 * level * width + j0;
#pragma unroll 4
        for (int rr = 0; rr < rows; rr++, krow += width) {
            T kv[4];
#pragma unroll
            for (int y = 0; y < 4; y++) kv[y] = col_ok[y] ? krow[y] : (T)0;
            const uint32_t packed = *reinterpret_cast<const uint32_t *>(dig + (size_t)rr * KS_BT + bg * 4);
#pragma unroll
            for (int x = 0; x < 4; x++) {
                const uint32_t d = (packed >> (8 * x)) & 0xFFu;
#pragma unroll
                for (int y = 0; y < 4; y++) acc[x][y] += (T)d * kv[y];
            }
        }
    }
    (void)vec_ok;
    // ---- epilogue: out = [body] - (acc - B/2 * colsum) ----
#pragma unroll
    for (int x = 0; x < 4; x++) {
        const int gb = b0 + bg * 4 + x;
        if (gb >= batch) continue;
#pragma unroll
        for (int y = 0; y < 4; y++) {
            const int j = j0 + y;
            if (j >= width) continue;
            T v = (T)half_b * colsum[j] - acc[x][y];
            if (j == body_col) v += in[(size_t)gb * in_stride + n_in];
            out[(size_t)gb * out_stride + j] = v;
        }
    }
}

template <typename T>
__global__ void ks_colsum_kernel(const T *__restrict__ ksk, T *__restrict__ colsum, size_t rows, int width) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= width) return;
    T s = 0;
    for (size_t r = 0; r < rows; r++) s += ksk[r * width + j];
    colsum[j] = s;
}

int ks_colsum_launch(const void *ksk, void *colsum, int n_in, int n_out, int level, int base_log, int bits,
                     cudaStream_t st) {
    (void)base_log;
    const int width = n_out + 1;
    const size_t rows = (size_t)n_in * level;
    if (bits == 64) ks_colsum_kernel<uint64_t><<<(width + 127) / 128, 128, 0, st>>>((const uint64_t *)ksk, (uint64_t *)colsum, rows, width);
    else ks_colsum_kernel<uint32_t><<<(width + 127) / 128, 128, 0, st>>>((const uint32_t *)ksk, (uint32_t *)colsum, rows, width);
    return (int)cudaGetLastError();
}

// ---- a handful of ciphertexts (the single-bootstrap latency path): split the key ROWS over the whole GPU -------------------------
// ks_kernel gives a batch of <= 32 ciphertexts to six CTAs (one per 128 output columns), each walking all n_in*l key rows: 0.12 ms for ONE
// ciphertext at PARAM_MESSAGE_2_CARRY_2, 6 % of a single KS+PBS.  Here a CTA owns a few (KS_SK_IC) input coefficients x 128 columns for all (<= 8)
// ciphertexts and adds its partial sums into the output with atomic adds: wrapping addition is associative and commutative, so the
// result is bit-identical to the serial loop whatever the order.  The output is pre-set to body + B/2 * colsum by ks_splitk_init_kernel.
constexpr int KS_SK_MAX = 8;       // ciphertexts

template <typename T>
__global__ void ks_splitk_init_kernel(const T *__restrict__ in, size_t in_stride, T *__restrict__ out, size_t out_stride,
                                      const T *__restrict__ colsum, int n_in, int width, int base_log, int batch, int body_col) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (j >= width || b >= batch) return;
    T v = (T)(1u << (base_log - 1)) * colsum[j];
    if (j == body_col) v += in[(size_t)b * in_stride + n_in];
    out[(size_t)b * out_stride + j] = v;
}

template <typename T, int KS_SK_IC, int UNR>
__global__ void __launch_bounds__(128)
ks_splitk_kernel(const T *__restrict__ in, size_t in_stride, T *__restrict__ out, size_t out_stride, const T *__restrict__ ksk,
                 int n_in, int width, int level, int base_log, int batch) {
    __shared__ __align__(8) unsigned char dig[KS_SK_IC * 8 * KS_SK_MAX];     // [(ii*level + t)][b], level <= 8
    const int tid = threadIdx.x, i0 = blockIdx.x * KS_SK_IC, j = blockIdx.y * 128 + tid;
    const uint32_t half_b = 1u << (base_log - 1);
    const int ni = min(KS_SK_IC, n_in - i0);
    for (int e = tid; e < ni * batch; e += 128) {
        const int ii = e / batch, b = e % batch;
        T state = decomp_init_state<T>(in[(size_t)b * in_stride + i0 + ii], base_log, level);
        for (int t = 0; t < level; t++) dig[(ii * level + t) * KS_SK_MAX + b] = (unsigned char)((int)decomp_next_digit<T>(state, base_log) + (int)half_b);
    }
    __syncthreads();
    if (j >= width) return;
    T acc[KS_SK_MAX];
#pragma unroll
    for (int b = 0; b < KS_SK_MAX; b++) acc[b] = 0;
    const T *krow = ksk + (size_t)i0 * level * width + j;
    const int rows = ni * level;
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
#pragma unroll UNR
    for (int rr = 0; rr < rows; rr++, krow += width) {
        // the key is read once: evict-first, so that streaming its 61 MB does not push the bootstrap key of the bootstrap that follows out of L2
        T kv;
        if constexpr (sizeof(T) == 8) asm volatile("ld.global.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(kv) : "l"(krow), "l"(pol));
        else asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(kv) : "l"(krow), "l"(pol));
        const uint2 pk = *reinterpret_cast<const uint2 *>(dig + rr * KS_SK_MAX);
#pragma unroll
        for (int b = 0; b < KS_SK_MAX; b++) acc[b] += (T)(((b < 4 ? pk.x : pk.y) >> (8 * (b & 3))) & 0xFFu) * kv;
    }
#pragma unroll
    for (int b = 0; b < KS_SK_MAX; b++) {
        if (b >= batch) break;
        T *o = out + (size_t)b * out_stride + j;
        if constexpr (sizeof(T) == 8) atomicAdd(reinterpret_cast<unsigned long long *>(o), (unsigned long long)(0 - acc[b]));
        else atomicAdd(reinterpret_cast<unsigned int *>(o), (unsigned int)(0 - acc[b]));
    }
}

template <typename T> static int ks_splitk_launch(const KSLaunch &a, cudaStream_t st) {
    const int width = a.n_out + 1, body = a.body_col < 0 ? a.n_out : a.body_col;
    ks_splitk_init_kernel<T><<<dim3((width + 127) / 128, a.batch), 128, 0, st>>>((const T *)a.in, a.in_stride, (T *)a.out, a.out_stride, (const T *)a.colsum,
                                                                               a.n_in, width, a.base_log, a.batch, body);
    // 4 input coefficients (20 key rows at l_ks = 5) per CTA, every load of a thread in flight at once: 3072 CTAs at PARAM_MESSAGE_2_CARRY_2.
    // Measured for one ciphertext (keyswitch alone / KS+PBS, us): 16 coefficients, unroll 4: 45.6 / 1756; 8, 8: 40.4 / 1750; 4, 20: 35.0 / 1738;
    // 3, 15: 33.2 / 1738; 2, 10: 41.1 / 1749; 1, 8: 59.6 / 1788.
    constexpr int IC = 4, UNR = 20;
    ks_splitk_kernel<T, IC, UNR><<<dim3((a.n_in + IC - 1) / IC, (width + 127) / 128), 128, 0, st>>>((const T *)a.in, a.in_stride, (T *)a.out, a.out_stride,
                                                                                                (const T *)a.ksk, a.n_in, width, a.level, a.base_log, a.batch);
    return (int)cudaGetLastError();
}

int ks_launch(const KSLaunch &a, cudaStream_t st) {
    if (a.batch <= 0) return 0;
    if (a.base_log < 1 || a.base_log > 7 || a.level < 1 || a.level * a.base_log >= a.bits) return -1;
    if (a.batch <= KS_SK_MAX && a.level <= 8 && a.sub <= 1 && !a.no_body) a.launches = 2;
    if (a.batch <= KS_SK_MAX && a.level <= 8 && a.sub <= 1 && !a.no_body) return a.bits == 64 ? ks_splitk_launch<uint64_t>(a, st) : ks_splitk_launch<uint32_t>(a, st);
    dim3 grid((a.batch + KS_BT - 1) / KS_BT, (a.n_out + 1 + KS_JT - 1) / KS_JT);
    size_t smem = (size_t)KS_IC * a.level * KS_BT;
    if (a.bits == 64)
        ks_kernel<uint64_t><<<grid, KS_THREADS, smem, st>>>((const uint64_t *)a.in, a.in_stride, (uint64_t *)a.out, a.out_stride,
                                                            (const uint64_t *)a.ksk, (const uint64_t *)a.colsum, a.n_in, a.n_out,
                                                            a.level, a.base_log, a.batch, a.body_col < 0 ? a.n_out : a.body_col);
    else
        ks_kernel<uint32_t><<<grid, KS_THREADS, smem, st>>>((const uint32_t *)a.in, a.in_stride, (uint32_t *)a.out, a.out_stride,
                                                            (const uint32_t *)a.ksk, (const uint32_t *)a.colsum, a.n_in, a.n_out,
                                                            a.level, a.base_log, a.batch, a.body_col < 0 ? a.n_out : a.body_col);
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

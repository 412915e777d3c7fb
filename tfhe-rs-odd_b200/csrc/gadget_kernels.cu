// gadget_kernels.cu -- the integer kernels around the blind rotation that the fork's gadget engine adds.
//
// Reference (tfhe/src/gadget/engine/bootstrapping.rs):
//   mvb_bootstrap :567-622                 one blind rotation of a common-factor accumulator (entry a14), then for every
//                                          function f_i:  GLWE x v_i (polynomial_karatsuba_wrapping_mul, a plaintext polynomial
//                                          with p non-zero coefficients, create_vi_for_mvb :503-543) and sample extraction
//   pack_into_new_accumulator :690-773     p LWE ciphertexts -> one GLWE accumulator: re-order / negate, packing keyswitch of
//                                          each (algorithms/lwe_packing_keyswitch.rs:100-187), then sum of monomial-shifted copies
// All of it is wrapping integer arithmetic: results are bit-identical to the reference loops.
#include "kernels.h"

namespace b200tfhe {

// ---- out[b][i] = sample_extract_0( glwe[b] * poly[i] )  -----------------------------------------------------------------
// grid (batch, n_polys); the plaintext polynomial is compacted to its non-zero terms in shared memory (the MVB polynomials
// have p of them), one mask polynomial at a time is staged in shared memory.
// (A * v)[c] = sum_s v[s] * A[c - s]  with  A[c - s] := -A[N + c - s] for c < s  (X^N = -1).
template <typename T>
__global__ void __launch_bounds__(256)
glwe_mul_plain_extract_kernel(const T *__restrict__ glwe, const T *__restrict__ polys, T *__restrict__ out, int k, int N) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    T *A = reinterpret_cast<T *>(sm_raw);                 // N
    T *val = A + N;                                       // N  (non-zero coefficients)
    int *pos = reinterpret_cast<int *>(val + N);          // N
    __shared__ int nnz;
    __shared__ unsigned long long body_acc;
    const int b = blockIdx.x, i = blockIdx.y, tid = threadIdx.x;
    const T *v = polys + (size_t)i * N;
    const T *g = glwe + (size_t)b * (k + 1) * N;
    T *o = out + ((size_t)b * gridDim.y + i) * ((size_t)k * N + 1);
    if (tid == 0) { nnz = 0; body_acc = 0; }
    __syncthreads();
    for (int s = tid; s < N; s += blockDim.x) {
        const T x = v[s];
        if (x != 0) { const int q = atomicAdd(&nnz, 1); pos[q] = s; val[q] = x; }
    }
    __syncthreads();
    const int cnt = nnz;
    for (int j = 0; j < k; j++) {
        for (int c = tid; c < N; c += blockDim.x) A[c] = g[(size_t)j * N + c];
        __syncthreads();
        // extracted mask: out[j*N + t] = P[0] (t = 0), -P[N - t] otherwise   (glwe_sample_extraction.rs:91-147)
        for (int t = tid; t < N; t += blockDim.x) {
            const int c = (t == 0) ? 0 : N - t;
            T acc = 0;
            for (int q = 0; q < cnt; q++) {
                const int d = c - pos[q];
                const T a = A[d & (N - 1)];
                acc += val[q] * (d < 0 ? (T)(0 - a) : a);
            }
            o[(size_t)j * N + t] = (t == 0) ? acc : (T)(0 - acc);
        }
        __syncthreads();
    }
    // body: coefficient 0 of B * v
    {
        const T *B = g + (size_t)k * N;
        T acc = 0;
        for (int q = tid; q < cnt; q += blockDim.x) {
            const int s = pos[q];
            acc += val[q] * (s == 0 ? B[0] : (T)(0 - B[N - s]));
        }
        atomicAdd(&body_acc, (unsigned long long)acc);
        __syncthreads();
        if (tid == 0) o[(size_t)k * N] = (T)body_acc;
    }
}

int glwe_mul_plain_extract_launch(const void *glwe, const void *polys, void *out, int k, int N, int n_polys, int batch, int bits,
                                  cudaStream_t st) {
    if (batch <= 0 || n_polys <= 0) return 0;
    dim3 grid(batch, n_polys);
    if (bits == 64) {
        size_t smem = (size_t)N * (2 * sizeof(uint64_t) + sizeof(int));
        if (cudaError_t e = ensure_dynamic_smem((const void *)glwe_mul_plain_extract_kernel<uint64_t>, smem)) return (int)e;
        glwe_mul_plain_extract_kernel<uint64_t><<<grid, 256, smem, st>>>((const uint64_t *)glwe, (const uint64_t *)polys, (uint64_t *)out, k, N);
    } else {
        size_t smem = (size_t)N * (2 * sizeof(uint32_t) + sizeof(int));
        if (cudaError_t e = ensure_dynamic_smem((const void *)glwe_mul_plain_extract_kernel<uint32_t>, smem)) return (int)e;
        glwe_mul_plain_extract_kernel<uint32_t><<<grid, 256, smem, st>>>((const uint32_t *)glwe, (const uint32_t *)polys, (uint32_t *)out, k, N);
    }
    return (int)cudaGetLastError();
}

// ---- pack_into_new_accumulator, step 1: the p "accumulator elements" of every accumulator (:699-716) --------------------
// in: [batch][count][len] LWE ciphertexts; out: [batch][p][len]:
//   slot kk even: in[kk/2]; slot kk odd: -in[(p+1)/2 + (kk-1)/2]; a missing input is the zero ciphertext.
template <typename T>
__global__ void pack_elements_kernel(const T *__restrict__ in, T *__restrict__ out, int count, int p, size_t len, size_t total) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const size_t c = e % len, slot = e / len;
    const int kk = (int)(slot % p);
    const size_t b = slot / p;
    const int src = (kk % 2 == 0) ? kk / 2 : (p + 1) / 2 + (kk - 1) / 2;
    T v = 0;
    if (src < count) {
        v = in[(b * count + src) * len + c];
        if (kk % 2) v = (T)(0 - v);
    }
    out[e] = v;
}

// ---- step 3: out = sum_kk sum_{i < rep_kk} G_kk * X^{off_kk + i}  -  sum_{i < size/2} G_0 * X^{N - size/2 + i}   (:733-766)
// size = N / p; rep_0 = size/2, off_0 = 0; rep_kk = size, off_kk = size/2 + (kk-1)*size.  G: [batch][p][(k+1)N].
// (G * X^d)[c] = G[c - d] for c >= d, -G[N + c - d] otherwise (polynomial_wrapping_monic_monomial_mul_assign, d < N).
template <typename T>
__global__ void __launch_bounds__(256)
pack_rotate_sum_kernel(const T *__restrict__ G, T *__restrict__ out, int p, int k1, int N) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    T *P = reinterpret_cast<T *>(sm_raw);     // one polynomial of one element
    const int b = blockIdx.x, j = blockIdx.y, tid = threadIdx.x;
    const int size = N / p, half = size / 2;
    constexpr int MAXC = 32;                  // coefficients per thread (N <= 8192)
    T acc[MAXC];
#pragma unroll
    for (int u = 0; u < MAXC; u++) acc[u] = 0;
    for (int kk = 0; kk < p; kk++) {
        const T *src = G + (((size_t)b * p + kk) * k1 + j) * N;
        __syncthreads();
        for (int c = tid; c < N; c += blockDim.x) P[c] = src[c];
        __syncthreads();
        const int rep = kk == 0 ? half : size, off = kk == 0 ? 0 : half + (kk - 1) * size;
#pragma unroll
        for (int u = 0; u < MAXC; u++) {
            const int c = tid + u * 256;
            if (c >= N) break;
            T s = 0;
            for (int i = 0; i < rep; i++) {
                const int d = c - off - i;
                const T a = P[d & (N - 1)];
                s += (d < 0) ? (T)(0 - a) : a;
            }
            if (kk == 0) {
                for (int i = 0; i < half; i++) {   // the "last one": -G_0 * X^{N - half + i}
                    const int d = c - (N - half + i);
                    const T a = P[d & (N - 1)];
                    s -= (d < 0) ? (T)(0 - a) : a;
                }
            }
            acc[u] += s;
        }
    }
#pragma unroll
    for (int u = 0; u < MAXC; u++) {
        const int c = tid + u * 256;
        if (c < N) out[((size_t)b * k1 + j) * N + c] = acc[u];
    }
}

int pack_elements_launch(const void *in, void *out, int count, int p, size_t len, int batch, int bits, cudaStream_t st) {
    const size_t total = (size_t)batch * p * len;
    if (total == 0) return 0;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64) pack_elements_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)in, (uint64_t *)out, count, p, len, total);
    else pack_elements_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)in, (uint32_t *)out, count, p, len, total);
    return (int)cudaGetLastError();
}

int pack_rotate_sum_launch(const void *G, void *out, int p, int k, int N, int batch, int bits, cudaStream_t st) {
    if (batch <= 0) return 0;
    if (N > 32 * 256) return -1;
    dim3 grid(batch, k + 1);
    size_t smem = (size_t)N * (bits / 8);
    if (bits == 64) {
        if (cudaError_t e = ensure_dynamic_smem((const void *)pack_rotate_sum_kernel<uint64_t>, smem)) return (int)e;
        pack_rotate_sum_kernel<uint64_t><<<grid, 256, smem, st>>>((const uint64_t *)G, (uint64_t *)out, p, k + 1, N);
    } else {
        if (cudaError_t e = ensure_dynamic_smem((const void *)pack_rotate_sum_kernel<uint32_t>, smem)) return (int)e;
        pack_rotate_sum_kernel<uint32_t><<<grid, 256, smem, st>>>((const uint32_t *)G, (uint32_t *)out, p, k + 1, N);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

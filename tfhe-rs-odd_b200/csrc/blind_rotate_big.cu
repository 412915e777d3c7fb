// blind_rotate_big.cu -- blind rotation / PBS for N = 8192 (PARAM_MESSAGE_3_CARRY_3), one CTA per ciphertext.
// Stage functions and the data-flow rationale: br_big.cuh.  Same entry semantics as blind_rotate.cu.
#include "br_big.cuh"
#include "kernels.h"

namespace b200tfhe {

template <typename T, int K, int L, int BL>
__global__ void __launch_bounds__(256, 1)
blind_rotate_big_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                        const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                        size_t out_stride, int n, int batch, int mode, const BigTables *__restrict__ tb) {
    typedef BigCore<T, K, L, BL> C;
    constexpr int K1 = K + 1, N = C::N, P = C::P;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *acc = reinterpret_cast<T *>(smem_raw);
    cplx *S1 = reinterpret_cast<cplx *>(smem_raw + C::ACC_BYTES);
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + C::ACC_BYTES + C::XCH_BYTES);
    const int tid = threadIdx.x;
    const int ct = blockIdx.x;
    if (ct >= batch) return;

    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = tid; i <= n; i += 256) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    __syncthreads();
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int r = 0; r < K1; r++)
            for (int c = tid; c < N; c += 256) acc[(size_t)r * N + c] = rot_read<T, N>(lut + (size_t)r * N, c, a0);
    }
    __syncthreads();

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw = bsk + (size_t)i * C::GGSW_CPLX;
        cplx F[K1][P];
        static_for<0, K1>([&](auto j) { static_for<0, P>([&](auto p) { F[decltype(j)::value][decltype(p)::value] = cplx{0.0, 0.0}; }); });
        // level L down to 1 (t ascending), rows ascending: the reference's summation order (ggsw.rs:524-577)
        static_for<0, L>([&](auto t) {
            for (int r = 0; r < K1; r++) {
                cplx x[P];
                C::template digits<decltype(t)::value>(tid, acc + (size_t)r * N, a, x);
                __syncthreads();               // previous user of S1 is done
                C::fwdA(tid, x, S1, tb);
                __syncthreads();
                C::fwdB_load(tid, x, S1);
                __syncwarp();
                C::fwdB_pass1(tid, x, S1, tb);
                __syncwarp();
                C::fwdB_pass2(tid, x, S1);
                C::mac(tid, L - 1 - decltype(t)::value, r, ggsw, x, F);
            }
        });
        static_for<0, K1>([&](auto j) {
            cplx x[P];
            static_for<0, P>([&](auto pp) { x[decltype(pp)::value] = F[decltype(j)::value][brev(decltype(pp)::value, C::LOGP)]; });
            __syncthreads();
            C::invA(tid, x, S1, tb);
            __syncwarp();
            C::invB_load(tid, x, S1);
            __syncwarp();
            C::invB_store(tid, x, S1);
            __syncthreads();
            C::invC(tid, x, S1, tb);
            C::accumulate(tid, x, acc + (size_t)decltype(j)::value * N);
        });
        __syncthreads();                        // acc complete before the next iteration's rotated reads
    }

    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = tid; c < K1 * N; c += 256) o[c] = acc[c];
    } else {
        for (int e = tid; e < K * N; e += 256) {
            const int r = e / N, t = e % N;
            o[e] = (t == 0) ? acc[(size_t)r * N] : (T)(0 - acc[(size_t)r * N + (N - t)]);
        }
        if (tid == 0) o[(size_t)K * N] = acc[(size_t)K * N];
    }
}

template <typename T>
__global__ void __launch_bounds__(256) bsk_convert_big_kernel(const T *__restrict__ std_bsk, cplx *__restrict__ out, size_t npoly,
                                                               const BigTables *__restrict__ tb) {
    typedef BigCore<T, 1, 1, 1> C;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx *S1 = reinterpret_cast<cplx *>(smem_raw);
    const int tid = threadIdx.x;
    for (size_t pi = blockIdx.x; pi < npoly; pi += gridDim.x) {
        cplx x[C::P];
        C::key_input(tid, std_bsk + pi * C::N, x);
        __syncthreads();
        C::fwdA(tid, x, S1, tb);
        __syncthreads();
        C::fwdB_load(tid, x, S1);
        __syncwarp();
        C::fwdB_pass1(tid, x, S1, tb);
        __syncwarp();
        C::fwdB_pass2(tid, x, S1);
        cplx *o = out + pi * C::M + tid;
        static_for<0, C::P>([&](auto p) { o[decltype(p)::value * 256] = x[decltype(p)::value]; });
    }
}

template <typename T, int K, int L, int BL> static int launch_big(const BRLaunch &a, cudaStream_t st) {
    typedef BigCore<T, K, L, BL> C;
    auto kern = blind_rotate_big_kernel<T, K, L, BL>;
    size_t smem = C::ACC_BYTES + C::XCH_BYTES + ((((size_t)a.n + 1) * 2 + 15) & ~(size_t)15);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    kern<<<a.batch, 256, smem, st>>>((const T *)a.lwe_in, a.lwe_stride, (const T *)a.luts, a.lut_idx, (const cplx *)a.bsk,
                                     (T *)a.out, a.out_stride, a.n, a.batch, a.mode, (const BigTables *)a.tw_fwd);
    return (int)cudaGetLastError();
}

#define BIG_INSTANCES(X) \
    X(uint64_t, 1, 2, 15)

int br_big_supported(const Params &p) {
#define X(T, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 8192 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return 256;
    BIG_INSTANCES(X)
#undef X
    return 0;
}
int br_big_launch(const Params &p, const BRLaunch &a, cudaStream_t st) {
#define X(T, K, L, BL) \
    if (p.bits == (int)sizeof(T) * 8 && p.N == 8192 && p.k == K && p.pbs_level == L && p.pbs_base_log == BL) return launch_big<T, K, L, BL>(a, st);
    BIG_INSTANCES(X)
#undef X
    return -1;
}
int bsk_convert_big_launch(const Params &p, const void *std_bsk, void *out, const void *tables, cudaStream_t st) {
    if (p.bits != 64) return -1;
    size_t npoly = (size_t)p.n * p.pbs_level * (p.k + 1) * (p.k + 1);
    auto kern = bsk_convert_big_kernel<uint64_t>;
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, 65536)) return (int)e;
    int grid = (int)(npoly < 148 * 3 ? npoly : 148 * 3);
    kern<<<grid, 256, 65536, st>>>((const uint64_t *)std_bsk, (cplx *)out, npoly, (const BigTables *)tables);
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

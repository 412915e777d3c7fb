// blind_rotate_big.cu -- blind rotation / PBS for N = 8192 (PARAM_MESSAGE_3_CARRY_3), one CTA per ciphertext.
// Stage functions and the data-flow rationale: br_big.cuh.  Same entry semantics as blind_rotate.cu.
#include "br_big.cuh"
#include "tmem_util.cuh"
#include "kernels.h"

namespace b200tfhe {

#if defined(__CUDA_ARCH__)
DEV void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
                 "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
                 "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}
DEV void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr)
                 : "memory");
}
#endif

template <typename T, int K, int L, int BL>
__global__ void __launch_bounds__(256, 1)
blind_rotate_big_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                        const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                        size_t out_stride, int n, int batch, int mode, const BigTables *__restrict__ tb) {
#if defined(__CUDA_ARCH__)     // device pass only: the tcgen05 helpers of tmem_util.cuh have no host version
    typedef BigCore<T, K, L, BL> C;
    constexpr int K1 = K + 1, N = C::N, P = C::P;
    constexpr size_t TW_BYTES = 8 * 256 * sizeof(cplx);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *acc = reinterpret_cast<T *>(smem_raw);
    cplx *S1 = reinterpret_cast<cplx *>(smem_raw + C::ACC_BYTES);
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + C::ACC_BYTES + C::XCH_BYTES);          // ta | tq, 2 x [4][256]
    uint16_t *ahat = reinterpret_cast<uint16_t *>(smem_raw + C::ACC_BYTES + C::XCH_BYTES + TW_BYTES);
    __shared__ uint32_t tmem_base_smem;
    const int tid = threadIdx.x;
    const int ct = blockIdx.x;
    if (ct >= batch) return;
    if (tid < 32) {   // tensor-memory columns [0,256): the Fourier accumulators of the 8 warps; [256,320): second-level digits
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "n"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tmem_fence_before();
    for (int e = tid; e < 8 * 256; e += 256) tw_s[e] = (&tb->ta[0][0])[e];                  // ta and tq are contiguous
    const cplx(*ta)[256] = reinterpret_cast<const cplx(*)[256]>(tw_s);
    const cplx(*tq)[256] = reinterpret_cast<const cplx(*)[256]>(tw_s + 4 * 256);

    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int i = tid; i <= n; i += 256) ahat[i] = (uint16_t)(modswitch<T>(lwe[i], C::LOGN) & (2 * N - 1));
    __syncthreads();
    tmem_fence_after();
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N;
        const int a0 = (2 * N - (int)ahat[n]) & (2 * N - 1);
        for (int r = 0; r < K1; r++)
            for (int c = tid; c < N; c += 256) acc[(size_t)r * N + c] = rot_read<T, N>(lut + (size_t)r * N, c, a0);
    }
    __syncthreads();

    // Fourier accumulators F[j][p] live in TENSOR MEMORY between the (level, row) steps of an iteration (tcgen05.st/ld, 32x32b:
    // thread = TMEM lane; warps w and w+4 share a lane quadrant and use columns [0,128) / [128,256)).  That frees 128 registers,
    // which hold the step's 32 bootstrap-key values instead: their loads are issued BEFORE the step's digits + forward FFT, so
    // the L2 latency that used to be exposed at every multiply-accumulate (long_scoreboard 4.2 / issue,
    // profiles/r01_ncu_other_kernels.json) is covered by the transform.
    const int warp = tid >> 5;
    const uint32_t t_f = tmem_base_smem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 128);
    // L = 2: the decomposition of a row runs ONCE per iteration; the level-(L-1) digits wait in tensor memory (16 packed words per
    // thread and row) for their own step, which keeps the reference's summation order (level-major) at half the integer work
    const uint32_t t_dig = tmem_base_smem + ((uint32_t)((warp & 3) * 32) << 16) + 256u + (uint32_t)((warp >> 2) * (16 * K1));

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw = bsk + (size_t)i * C::GGSW_CPLX;
        // level L down to 1 (t ascending), rows ascending: the reference's summation order (ggsw.rs:524-577); both loops
        // rolled: one copy of the forward transform in the instruction stream
#pragma unroll 1
        for (int s = 0; s < L * K1; s++) {
            const int t = s / K1, r = s % K1, lvl = L - 1 - t;
            // key values of output polynomial 0: issued before the transform; those of polynomial j+1 while polynomial j is
            // multiplied (64 registers in flight at a time: 128 spilled)
            const cplx *gbase = ggsw + (size_t)(lvl * K1 + r) * K1 * C::M + tid;
            cplx gv[P];
            static_for<0, P>([&](auto p) { gv[decltype(p)::value] = ldg_stream(gbase + decltype(p)::value * 256); });
            cplx x[P];
            if constexpr (L == 2) {
                uint32_t packed[P];
                if (t == 0) {
                    C::digits_both(tid, acc + (size_t)r * N, a, x, packed);
                    tmem_st16(t_dig + (uint32_t)(r * 16), packed);     // completed by the tmem_wait_st at the end of the step
                } else {
                    tmem_ld16(t_dig + (uint32_t)(r * 16), packed);
                    tmem_wait_ld();
                    C::digits_unpack(packed, x);
                }
            } else {
                C::digits_rt(tid, acc + (size_t)r * N, a, t, x);
            }
            __syncthreads();               // previous user of S1 is done
            C::fwdA(tid, x, S1, ta, tq);
            __syncthreads();
            C::fwdB_load(tid, x, S1);
            __syncwarp();
            C::fwdB_pass1(tid, x, S1, tb);
            __syncwarp();
            C::fwdB_pass2(tid, x, S1);
            static_for<0, K1>([&](auto j) {
                cplx f[P], gn[P];
                if constexpr (decltype(j)::value + 1 < K1) {
                    static_for<0, P>([&](auto p) { gn[decltype(p)::value] = ldg_stream(gbase + (size_t)(decltype(j)::value + 1) * C::M + decltype(p)::value * 256); });
                }
                if (s == 0) {
                    static_for<0, P>([&](auto p) { f[decltype(p)::value] = cplx{0.0, 0.0}; });
                } else {
                    uint32_t v0[32], v1[32];
                    tmem_ld32(t_f + decltype(j)::value * 64, v0);
                    tmem_ld32(t_f + decltype(j)::value * 64 + 32, v1);
                    tmem_wait_ld();
                    static_for<0, 8>([&](auto e) {
                        f[decltype(e)::value] = cplx_from_words(v0, decltype(e)::value);
                        f[8 + decltype(e)::value] = cplx_from_words(v1, decltype(e)::value);
                    });
                }
                C::mac_one_poly(gv, x, f);
                tmem_store_cplx8(t_f + decltype(j)::value * 64, f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7]);
                tmem_store_cplx8(t_f + decltype(j)::value * 64 + 32, f[8], f[9], f[10], f[11], f[12], f[13], f[14], f[15]);
                if constexpr (decltype(j)::value + 1 < K1) {
                    static_for<0, P>([&](auto p) { gv[decltype(p)::value] = gn[decltype(p)::value]; });
                }
            });
            tmem_wait_st();
        }
        static_for<0, K1>([&](auto j) {
            cplx x[P];
            {
                uint32_t v0[32], v1[32];
                tmem_ld32(t_f + decltype(j)::value * 64, v0);
                tmem_ld32(t_f + decltype(j)::value * 64 + 32, v1);
                tmem_wait_ld();
                static_for<0, P>([&](auto pp) {
                    constexpr int p = brev(decltype(pp)::value, C::LOGP);
                    if constexpr (p < 8) x[decltype(pp)::value] = cplx_from_words(v0, p);
                    else x[decltype(pp)::value] = cplx_from_words(v1, p - 8);
                });
            }
            __syncthreads();
            C::invA(tid, x, S1, tb);
            __syncwarp();
            C::invB_load(tid, x, S1);
            __syncwarp();
            C::invB_store(tid, x, S1);
            __syncthreads();
            C::invC(tid, x, S1, ta, tq);
            C::accumulate(tid, x, acc + (size_t)decltype(j)::value * N);
        });
        __syncthreads();                        // acc complete before the next iteration's rotated reads
    }

    tmem_fence_before();
    __syncthreads();
    if (tid < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "n"(512));
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
#endif
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
        C::fwdA(tid, x, S1, tb->ta, tb->tq);
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
    size_t smem = C::ACC_BYTES + C::XCH_BYTES + 8 * 256 * sizeof(cplx) + ((((size_t)a.n + 1) * 2 + 15) & ~(size_t)15);
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

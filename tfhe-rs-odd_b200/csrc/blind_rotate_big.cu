// blind_rotate_big.cu -- blind rotation / PBS for N = 8192 (PARAM_MESSAGE_3_CARRY_3), one CTA per ciphertext.
// Stage functions and the data-flow rationale: br_big.cuh.  Same entry semantics as blind_rotate.cu.
#include "br_big.cuh"
#include "tmem_util.cuh"
#include "kernels.h"

namespace b200tfhe {

// key loads of the product phase: every value is used once per SM and iteration (the wave shares the key through L2), so they do not
// allocate in the 32 KB of L1 that are left (285.4 -> 283.0 ms per 2048 ciphertexts)
__device__ __forceinline__ cplx ldg_noalloc(const cplx *p) {
    cplx v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.re), "=d"(v.im) : "l"(p));
    return v;
}


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
    if (tid < 32) {   // all 512 tensor-memory columns: the transformed polynomials of an iteration (see the main loop)
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

    // Round 2: ALL forward transforms of an iteration first, then the Fourier product.  The L*K1 transformed polynomials D[s]
    // (s = t*K1 + r, 16 complex values per thread each) are parked in TENSOR MEMORY (tcgen05.st/ld, 32x32b: thread = TMEM lane;
    // warps w and w+4 share a lane quadrant and use columns [0,256) / [256,512): 4 x 64 columns per thread = the whole TMEM).  A
    // thread reads back exactly what it stored (position p*256 + tid of every polynomial is produced and consumed by thread tid), so
    // no barrier separates the two phases.  What that buys over round 1 (Fourier accumulators in tensor memory, one
    // multiply-accumulate per transform): the forward phase holds no bootstrap-key registers (round 1: 128 of 255, 340 bytes of
    // spill), and the product phase is a software pipeline -- the key values and the D values of step s+1 are requested before step
    // s is multiplied -- with the Fourier accumulator of ONE output polynomial in registers, handed to the inverse transform directly.
    // Summation order: level L down to 1, rows ascending, as the reference (ggsw.rs:524-577).
    const int warp = tid >> 5;
    const uint32_t t_d = tmem_base_smem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 256);
    constexpr int NS = L * K1;     // transforms per iteration
    static_assert(NS * 64 * 2 <= 512, "the transformed polynomials of an iteration must fit the tensor memory");

    for (int i = 0; i < n; i++) {
        const int a = ahat[i];
        const cplx *ggsw = bsk + (size_t)i * C::GGSW_CPLX;
        // ---- forward phase: row by row; the decomposition of a row runs once, its level-(L-1) digits wait packed in registers ------
        // (measured and rejected: computing the register-only first half of the NEXT transform between the exchange loads of the
        // current one and their use -- 374 ms against 306 ms for 2048 ciphertexts: two 16-point register sets do not schedule well)
#pragma unroll 1
        for (int r = 0; r < K1; r++) {
            cplx x[P];
            uint32_t packed[P];
            if constexpr (L == 2) C::digits_both(tid, acc + (size_t)r * N, a, x, packed);
#pragma unroll
            for (int t = 0; t < L; t++) {
                if constexpr (L == 2) {
                    if (t == 1) C::digits_unpack(packed, x);
                } else {
                    C::digits_rt(tid, acc + (size_t)r * N, a, t, x);
                }
                C::fwdA_compute(tid, x, ta, tq);   // the register-only half runs BEFORE the barrier: a warp that is ahead works instead of waiting
                __syncthreads();               // previous user of S1 is done
                C::fwdA_store(tid, x, S1);
                __syncthreads();
                C::fwdB_load(tid, x, S1);
                __syncwarp();
                C::fwdB_pass1(tid, x, S1, tb);
                __syncwarp();
                C::fwdB_pass2(tid, x, S1);
                const uint32_t slot = t_d + (uint32_t)((t * K1 + r) * 64);
                tmem_store_cplx8(slot, x[0], x[1], x[2], x[3], x[4], x[5], x[6], x[7]);
                tmem_store_cplx8(slot + 32, x[8], x[9], x[10], x[11], x[12], x[13], x[14], x[15]);
            }
        }
        tmem_wait_st();
        // ---- product + inverse phase, one output polynomial at a time ----------------------------------------------------------------
#pragma unroll 1
        for (int j = 0; j < K1; j++) {      // rolled: one copy of the product + inverse code (the unrolled forward pair already fills the instruction cache)
            cplx f[P];
            static_for<0, P>([&](auto p) { f[decltype(p)::value] = cplx{0.0, 0.0}; });
            // half-steps h = 2 s + e: positions p = 8 e .. 8 e + 7 of transform s; the loads of half-step h + 1 are in flight while h is used
            auto key_ptr = [&](int s) {
                const int t = s / K1, r = s % K1, lvl = L - 1 - t;
                return ggsw + ((size_t)(lvl * K1 + r) * K1 + j) * C::M + tid;
            };
            // key values: a ring of KD register slices (8 positions each), KD - 1 half-steps ahead of their use: with 256 threads that
            // is (KD - 1) x 32 KB in flight against ~700 cycles of L2 latency; D values: one half-step ahead (tensor memory is close)
            constexpr int KD = 2, NH = 2 * NS;
            cplx g[KD][8];
            uint32_t d[2][32];
            auto load_key = [&](auto hh) {
                constexpr int h = decltype(hh)::value;
                const cplx *kp = key_ptr(h / 2) + (h & 1) * 8 * 256;
                static_for<0, 8>([&](auto e) { g[h % KD][decltype(e)::value] = ldg_noalloc(kp + decltype(e)::value * 256); });
            };
            static_for<0, KD - 1>([&](auto hh) { load_key(hh); });
            tmem_ld32(t_d, d[0]);
            static_for<0, NH>([&](auto h_) {
                constexpr int h = decltype(h_)::value, e8 = (h & 1) * 8;
                if constexpr (h + KD - 1 < NH) load_key(std::integral_constant<int, h + KD - 1>{});
                tmem_wait_ld();                       // the D values of half-step h (requested one half-step ago)
                if constexpr (h + 1 < NH) tmem_ld32(t_d + (uint32_t)(((h + 1) / 2) * 64 + ((h + 1) & 1) * 32), d[(h + 1) & 1]);
                static_for<0, 8>([&](auto e) {
                    const cplx kv = g[h % KD][decltype(e)::value];
                    const cplx dv = cplx_from_words(d[h & 1], decltype(e)::value);
                    cplx &o = f[e8 + decltype(e)::value];
                    o.re = fma_(kv.re, dv.re, o.re);
                    o.re = fma_(-kv.im, dv.im, o.re);
                    o.im = fma_(kv.re, dv.im, o.im);
                    o.im = fma_(kv.im, dv.re, o.im);
                });
            });
            cplx x[P];
            static_for<0, P>([&](auto pp) { x[decltype(pp)::value] = f[brev(decltype(pp)::value, C::LOGP)]; });
            C::invA_compute(tid, x, tb);        // likewise: butterflies and team twiddles before the barrier, only the store after it
            __syncthreads();
            C::invA_store(tid, x, S1);
            __syncwarp();
            C::invB_load(tid, x, S1);
            __syncwarp();
            C::invB_store(tid, x, S1);
            __syncthreads();
            C::invC(tid, x, S1, ta, tq);
            C::accumulate(tid, x, acc + (size_t)j * N);
        }
        // No barrier here when k >= 1: row 0 was completed before the two barriers of the last inverse transform, and row r > 0 is read
        // only after the barriers of row 0's forward transforms of the next iteration.
        if constexpr (K1 < 2) __syncthreads();
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

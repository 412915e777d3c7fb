// blind_rotate_mb.cu -- batched MULTI-BIT blind rotation / programmable bootstrap for sm_100a.
//
// Replaces multi_bit_(deterministic_)blind_rotate_assign / multi_bit_(deterministic_)programmable_bootstrap_lwe_ciphertext
// (tfhe/src/core_crypto/algorithms/lwe_multi_bit_programmable_bootstrapping.rs:295-1224) over a ciphertext list.
// The reference spends thread_count CPU threads per ciphertext on preparing GGSW_sum for upcoming groups while one thread
// runs the external products; here a CTA owns CTS ciphertexts for all n/G iterations exactly like blind_rotate.cu, and the
// monomial-weighted sum over the 2^G key ciphertexts of a group is folded into the Fourier multiply-accumulate
// (br_multibit.cuh): nothing of size GGSW is ever written, and the sequential depth of a bootstrap drops from n to n/G.
#include "br_multibit.cuh"
#include "kernels.h"

namespace b200tfhe {

template <typename T, int R, int K, int L, int BL, int CTS, int G> struct MBConfig {
    typedef MBCore<T, R, K, L, BL, G> C;
    static constexpr int K1 = K + 1;
    static constexpr int THREADS = CTS * K1 * R;
    static_assert(THREADS % 32 == 0, "teams must fill whole warps");
    static constexpr size_t FBUF_BYTES = (size_t)K1 * L * C::REGION * sizeof(cplx);   // per ciphertext
    static constexpr size_t ACC_BYTES = (size_t)K1 * C::N * sizeof(T);
    static constexpr size_t TW_BYTES = sizeof(TwTables<R>) + R * sizeof(cplx);       // twiddle tables + R-th roots
    __host__ __device__ static size_t deg_bytes(int n) { return (((size_t)(n / G) * C::NMONO * sizeof(uint16_t)) + 15) & ~(size_t)15; }
    static size_t smem_bytes(int n) { return CTS * (FBUF_BYTES + ACC_BYTES + deg_bytes(n)) + TW_BYTES; }
};

template <typename T, int R, int K, int L, int BL, int CTS, int G>
__global__ void __launch_bounds__(CTS *(K + 1) * R)
blind_rotate_mb_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const T *__restrict__ luts,
                       const uint32_t *__restrict__ lut_idx, const cplx *__restrict__ bsk, T *__restrict__ out,
                       size_t out_stride, int n, int batch, int mode, const cplx *__restrict__ tw_fwd_,
                       const cplx *__restrict__ tw_inv_) {
    typedef MBConfig<T, R, K, L, BL, CTS, G> CFG;
    typedef typename CFG::C C;
    constexpr int K1 = K + 1, N = C::N, REGION = C::REGION, NMONO = C::NMONO;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int team = threadIdx.x / R, lane = threadIdx.x % R;
    const int ct_local = team / K1, r = team % K1;
    int ct = blockIdx.x * CTS + ct_local;
    const bool live = ct < batch;
    if (!live) ct = batch - 1;  // ghost ciphertext: same work, no output (keeps barriers uniform)

    const int ngroups = n / G;
    const size_t deg_bytes = CFG::deg_bytes(n);
    cplx *fbuf_all = reinterpret_cast<cplx *>(smem_raw);
    T *acc_all = reinterpret_cast<T *>(smem_raw + CTS * CFG::FBUF_BYTES);
    unsigned char *deg_all = smem_raw + CTS * (CFG::FBUF_BYTES + CFG::ACC_BYTES);
    cplx *fbuf_ct = fbuf_all + (size_t)ct_local * K1 * L * REGION;
    cplx *fbuf_r = fbuf_ct + (size_t)r * L * REGION;
    T *acc_r = acc_all + (size_t)ct_local * K1 * N + (size_t)r * N;
    uint16_t *deg = reinterpret_cast<uint16_t *>(deg_all + ct_local * deg_bytes);

    cplx *tw_s = reinterpret_cast<cplx *>(deg_all + CTS * deg_bytes);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int e = threadIdx.x; e < TW_HALF; e += CFG::THREADS) {
        tw_s[e] = tw_fwd_[e];
        tw_s[TW_HALF + e] = tw_inv_[e];
    }
    cplx *wr = tw_s + 2 * TW_HALF;
    for (int m = threadIdx.x; m < R; m += CFG::THREADS) wr[m] = C::wr_entry(m);
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    // ---- prologue: monomial degrees of every group (modulus switch of the SUM of the selected mask elements,
    // lwe_multi_bit_programmable_bootstrapping.rs:47-64) and acc <- LUT * X^{-b_hat} (:375-392) ------------------
    const T *lwe = lwe_in + (size_t)ct * lwe_stride;
    for (int e = r * R + lane; e < ngroups * NMONO; e += K1 * R) {
        const int gi = e / NMONO, idx = e % NMONO + 1;
        deg[e] = C::group_degree(lwe + (size_t)gi * G, idx);
    }
    {
        const T *lut = luts + (size_t)(lut_idx ? lut_idx[ct] : 0) * K1 * N + (size_t)r * N;
        const int bhat = (int)(modswitch<T>(lwe[n], C::LOGN) & (2 * N - 1));
        const int a0 = (2 * N - bhat) & (2 * N - 1);
        for (int c = lane; c < N; c += R) acc_r[c] = rot_read<T, N>(lut, c, a0);
    }
    __syncthreads();

    auto ct_barrier = [&]() {
        if constexpr (CTS > 1 && (K1 * R) % 32 == 0 && CTS <= 15) {
            asm volatile("bar.sync %0, %1;" ::"r"(ct_local + 1), "n"(K1 * R) : "memory");
        } else {
            __syncthreads();
        }
    };

    // ---- n/G external products -------------------------------------------------------------------------------
    for (int gi = 0; gi < ngroups; gi++) {
        if constexpr (CTS > 1) __syncthreads();   // keep the CTA's ciphertexts on the same key group (L1 reuse, I-cache)
        C::digits_plain_fwd_pass1(lane, acc_r, fbuf_r, twf);
        __syncwarp();
#pragma unroll 1
        for (int t = 0; t < L; t++) {
            cplx x[R];
            C::fwd_pass2_load(lane, x, fbuf_r + t * REGION);
            __syncwarp();
            C::fwd_pass2_store(lane, x, fbuf_r + t * REGION);
        }
        ct_barrier();
        cplx F[R];
        C::mb_mac(lane, r, bsk + (size_t)gi * C::GROUP_CPLX, fbuf_ct, deg + gi * NMONO, wr, F);
        ct_barrier();
        C::inv_pass1(lane, F, fbuf_r, twi);
        __syncwarp();
        C::inv_pass2_store(lane, fbuf_r, acc_r);
        __syncwarp();
    }

    // ---- epilogue ---------------------------------------------------------------------------------------------
    if (!live) return;
    T *o = out + (size_t)ct * out_stride;
    if (mode == BR_MODE_GLWE) {
        for (int c = lane; c < N; c += R) o[(size_t)r * N + c] = acc_r[c];
    } else {
        if (r < K) {
            for (int t = lane; t < N; t += R) o[(size_t)r * N + t] = (t == 0) ? acc_r[0] : (T)(0 - acc_r[N - t]);
        } else if (lane == 0) {
            o[(size_t)K * N] = acc_r[0];
        }
    }
}

// ---- two-kernel latency path ---------------------------------------------------------------------------------------
// With few ciphertexts the fused kernel above leaves most SMs idle while one CTA per ciphertext walks n/G iterations that each
// read 2^G key ciphertexts.  The reference spends its spare threads on exactly that part (prepare_multi_bit_ggsw_mem_optimized on
// worker threads, :295-546); here EVERY (ciphertext, group) pair gets a CTA that materialises
//     GGSW'_g = GGSW_0 + sum_{idx>=1} GGSW_idx (.) FFT(X^{d_idx})                                   (:18-85)
// in the device's Fourier order, all groups at once across the whole GPU, and the single-ciphertext kernel of
// blind_rotate_lat.cu then runs n/G plain external products against its ciphertext's list: the sequential depth of a bootstrap
// is n/G iterations of the classic latency kernel.
template <typename T, int R, int K, int L, int BL, int G>
__global__ void __launch_bounds__(256)
mb_key_sum_kernel(const T *__restrict__ lwe_in, size_t lwe_stride, const cplx *__restrict__ bsk, cplx *__restrict__ out, int n) {
    typedef MBCore<T, R, K, L, BL, G> C;
    constexpr int NG = C::NG, N = C::N, NMONO = C::NMONO;
    static_assert(256 % R == 0, "a thread keeps one lane of the team layout");
    __shared__ cplx wr[R];
    __shared__ uint16_t deg[NMONO];
    const int gi = blockIdx.x, ct = blockIdx.y, ngroups = n / G;
    if (threadIdx.x < R) wr[threadIdx.x] = C::wr_entry(threadIdx.x);
    if (threadIdx.x >= 64 && threadIdx.x < 64 + NMONO)
        deg[threadIdx.x - 64] = C::group_degree(lwe_in + (size_t)ct * lwe_stride + (size_t)gi * G, (int)threadIdx.x - 64 + 1);
    __syncthreads();
    const int lane = threadIdx.x % R;
    cplx base[NMONO];
    int dgs[NMONO];
    static_for<0, NMONO>([&](auto g) {
        dgs[decltype(g)::value] = deg[decltype(g)::value];
        base[decltype(g)::value] = unit_root_pi((dgs[decltype(g)::value] * (1 - 4 * lane)) & (2 * N - 1), N);
    });
    const cplx *group = bsk + (size_t)gi * C::GROUP_CPLX;
    cplx *o = out + ((size_t)ct * ngroups + gi) * C::GGSW_CPLX;
#pragma unroll 2
    for (int e = threadIdx.x; e < (int)C::GGSW_CPLX; e += 256) {
        const int pos = (e / R) % R;     // e = polynomial * M + pos * R + lane
        cplx s = ldg_stream(group + e);
        static_for<1, NG>([&](auto g) {
            const cplx w = wr[(dgs[decltype(g)::value - 1] * pos) & (R - 1)];
            C::mac_one(s, cmul(base[decltype(g)::value - 1], w), ldg_stream(group + (size_t) decltype(g)::value * C::GGSW_CPLX + e));
        });
        o[e] = s;
    }
}

template <typename T, int R, int K, int L, int BL, int G>
static int launch_mb_lat(const Params &p, const BRLaunch &a, cudaStream_t st) {
    typedef MBCore<T, R, K, L, BL, G> C;
    const int ngroups = a.n / G;
    mb_key_sum_kernel<T, R, K, L, BL, G><<<dim3(ngroups, a.batch), 256, 0, st>>>((const T *)a.lwe_in, a.lwe_stride, (const cplx *)a.bsk,
                                                                                  (cplx *)a.mb_scratch, a.n);
    if (cudaError_t e = cudaGetLastError()) return (int)e;
    BRLaunch b = a;
    b.bsk = a.mb_scratch;
    a.launches = 2;
    return br_lat_plain_launch(p, b, ngroups, (size_t)ngroups * C::GGSW_CPLX, st);
}

template <typename T, int R, int K, int L, int BL, int CTS, int G>
static int launch_mb(const BRLaunch &a, cudaStream_t st) {
    typedef MBConfig<T, R, K, L, BL, CTS, G> CFG;
    auto kern = blind_rotate_mb_kernel<T, R, K, L, BL, CTS, G>;
    size_t smem = CFG::smem_bytes(a.n);
    if (cudaError_t e = ensure_dynamic_smem((const void *)kern, smem)) return (int)e;
    int grid = (a.batch + CTS - 1) / CTS;
    kern<<<grid, CFG::THREADS, smem, st>>>((const T *)a.lwe_in, a.lwe_stride, (const T *)a.luts, a.lut_idx,
                                           (const cplx *)a.bsk, (T *)a.out, a.out_stride, a.n, a.batch, a.mode,
                                           (const cplx *)a.tw_fwd, (const cplx *)a.tw_inv);
    return (int)cudaGetLastError();
}

// (T, R, K, L, BASE_LOG, CTS, G): PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_{2,3}_KS_PBS
// (shortint/parameters/multi_bit.rs:115-132,173-190) and the toy sets of the test-suite
#define MB_INSTANCES(X)               \
    X(uint64_t, 32, 1, 1, 22, 3, 2)   \
    X(uint64_t, 32, 1, 1, 21, 3, 3)   \
    X(uint64_t, 16, 1, 2, 12, 1, 2)   \
    X(uint64_t, 16, 1, 2, 12, 1, 3)
// fewer ciphertexts than 3 per SM: one per CTA (the single-bootstrap latency configuration)
#define MB_SMALL_BATCH_INSTANCES(X)   \
    X(uint64_t, 32, 1, 1, 22, 1, 2)   \
    X(uint64_t, 32, 1, 1, 21, 1, 3)

#define MB_MATCH(T, R, K, L, BL, G) \
    (p.bits == (int)sizeof(T) * 8 && p.N == 2 * R * R && p.k == K && p.pbs_level == L && p.pbs_base_log == BL && grouping == G)

int br_mb_supported(const Params &p, int grouping) {
    if (grouping < 2 || p.n % grouping != 0) return 0;
#define X(T, R, K, L, BL, CTS, G) \
    if (MB_MATCH(T, R, K, L, BL, G)) return R;
    MB_INSTANCES(X)
#undef X
    return 0;
}

// the latency path exists where the single-ciphertext kernel does: N = 2048, k = 1, one level
#define MB_LAT_INSTANCES(X)       \
    X(uint64_t, 32, 1, 1, 22, 2)  \
    X(uint64_t, 32, 1, 1, 21, 3)

size_t br_mb_lat_scratch_bytes(const Params &p, int grouping, int batch) {
    if (!br_lat_plain_supported(p) || batch <= 0) return 0;
#define X(T, R, K, L, BL, G) \
    if (MB_MATCH(T, R, K, L, BL, G)) return (size_t)batch * (size_t)(p.n / G) * MBCore<T, R, K, L, BL, G>::GGSW_CPLX * sizeof(cplx);
    MB_LAT_INSTANCES(X)
#undef X
    return 0;
}

int br_mb_launch(const Params &p, int grouping, const BRLaunch &a, cudaStream_t st) {
    if (a.batch <= 0) return 0;
    if (a.mb_scratch && a.batch <= a.mb_lat_max && br_lat_plain_supported(p)) {
#define X(T, R, K, L, BL, G) \
    if (MB_MATCH(T, R, K, L, BL, G)) return launch_mb_lat<T, R, K, L, BL, G>(p, a, st);
        MB_LAT_INSTANCES(X)
#undef X
    }
    if (a.batch < 3 * 148) {
#define X(T, R, K, L, BL, CTS, G) \
    if (MB_MATCH(T, R, K, L, BL, G)) return launch_mb<T, R, K, L, BL, CTS, G>(a, st);
        MB_SMALL_BATCH_INSTANCES(X)
#undef X
    }
#define X(T, R, K, L, BL, CTS, G) \
    if (MB_MATCH(T, R, K, L, BL, G)) return launch_mb<T, R, K, L, BL, CTS, G>(a, st);
    MB_INSTANCES(X)
#undef X
    return -1;
}

}  // namespace b200tfhe

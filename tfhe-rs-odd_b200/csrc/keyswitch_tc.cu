// keyswitch_tc.cu -- batched LWE keyswitch on the 5th-generation tensor cores (tcgen05.mma kind::i8), bit-exact.
//
// Same operation as keyswitch.cu (keyswitch_lwe_ciphertext, tfhe/src/core_crypto/algorithms/lwe_keyswitch.rs:96-170):
//     out = (0,...,0,b_in) - sum_i sum_t digit_t(a_i) * KSK[i][t]          (mod 2^W)
// The digits are tiny signed integers (|d| <= 2^(base_log-1) <= 64) and a key word splits into W/8 unsigned bytes, so
//     sum_r d_r * K[r][j] = sum_p 2^(8p) * S_p[j],     S_p[j] = sum_r d_r * byte_p(K[r][j])
// and every S_p is an exact int8 x uint8 -> int32 dot product (|S_p| <= rows * 64 * 255 < 2^31 for every supported set).
// That is one integer GEMM  [batch x rows] (s8)  x  [rows x (n_out+1) * W/8] (u8)  ->  s32, which is what tcgen05.mma
// kind::i8 computes; the epilogue recombines the W/8 byte planes of a column with shifts (wrapping), so the result is
// bit-identical to the serial reference loop.
//
// Kernel anatomy (one 128 x 256 output tile per CTA, 6 warps):
//   warp 0 (one lane)  TMA producer: 128B-swizzled K-major tiles of the digit matrix (128 x 128 B) and of the byte-plane
//                      key (256 x 128 B) into a 4-stage shared-memory ring, mbarrier complete_tx
//   warp 1 (one lane)  MMA issuer: 4 x tcgen05.mma (M128 N256 K32) per stage into a 256-column TMEM accumulator;
//                      tcgen05.commit frees the stage / signals the epilogue
//   warps 2..5         epilogue: tcgen05.ld 32x32b.x32 (thread = ciphertext row), recombine byte planes, subtract from
//                      the body, store
// Two small helper kernels: ks_digits_kernel (decomposition -> s8 matrix, once per batch) and ks_byteplanes_kernel
// (key -> byte-plane matrix, once per key upload).
#include "kernels.h"
#include <cuda.h>
#include <cstdio>

namespace b200tfhe {

namespace {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 128, TC_STAGES = 4, TC_UMMA_K = 32;
constexpr int TC_A_BYTES = TC_BM * TC_BK, TC_B_BYTES = TC_BN * TC_BK, TC_STAGE_BYTES = TC_A_BYTES + TC_B_BYTES;
constexpr int TC_THREADS = 192;
constexpr size_t TC_SMEM = (size_t)TC_STAGES * TC_STAGE_BYTES + 1024 /* alignment slack */ + 128 /* barriers */;

#if defined(__CUDA_ARCH__)
DEV uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
DEV void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
DEV void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
DEV void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity)
        : "memory");
}
DEV void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
                 "l"(map), "r"(bar), "r"(c0), "r"(c1)
                 : "memory");
}
// K-major operand tile, 128-byte rows, 128B swizzle: 8-row atoms of 1024 B (SBO), LBO = 1 (unused), descriptor version 1
DEV uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
DEV void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
DEV void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
DEV void tmem_ld32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
#endif

// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D = S32, A = signed 8 bit, B = unsigned 8 bit,
// both K-major, N = 256, M = 128
constexpr uint32_t TC_IDESC = (2u << 4) | (1u << 7) | (0u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

template <typename T>
__global__ void __launch_bounds__(TC_THREADS, 1)
ks_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const T *__restrict__ in,
             size_t in_stride, T *__restrict__ out, size_t out_stride, int n_in, int n_out, int k_blocks, int batch, int body_col) {
#if defined(__CUDA_ARCH__)
    constexpr int PL = sizeof(T);                 // byte planes per key word
    constexpr int COEF = TC_BN / PL;              // output coefficients per tile
    extern __shared__ unsigned char smem_raw[];
    __shared__ uint32_t tmem_base_smem;
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // 128B swizzle atoms need 1024 B alignment
    const uint32_t bars = smem_base + TC_STAGES * TC_STAGE_BYTES;          // full[4] | empty[4] | tmem_full
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (TC_STAGES + s); };
    const uint32_t tmem_full_bar = bars + 8u * (2 * TC_STAGES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tile = blockIdx.x, m_tile = blockIdx.y;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < TC_STAGES; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "n"(TC_BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_acc = tmem_base_smem;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < k_blocks; kb++) {
                const int s = kb % TC_STAGES;
                if (kb >= TC_STAGES) mbar_wait(empty_bar(s), ((kb / TC_STAGES) - 1) & 1);
                mbar_expect_tx(full_bar(s), TC_STAGE_BYTES);
                const uint32_t a_dst = smem_base + s * TC_STAGE_BYTES, b_dst = a_dst + TC_A_BYTES;
                tma_load_2d(a_dst, &map_a, full_bar(s), kb * TC_BK, m_tile * TC_BM);
                tma_load_2d(b_dst, &map_b, full_bar(s), kb * TC_BK, n_tile * TC_BN);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            for (int kb = 0; kb < k_blocks; kb++) {
                const int s = kb % TC_STAGES;
                mbar_wait(full_bar(s), (kb / TC_STAGES) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_addr = smem_base + s * TC_STAGE_BYTES, b_addr = a_addr + TC_A_BYTES;
#pragma unroll
                for (int k = 0; k < TC_BK / TC_UMMA_K; k++) {
                    // advancing K inside the 128-byte swizzle row: +32 bytes on the (unswizzled) start address
                    umma_i8(tmem_acc, umma_desc_sw128(a_addr + k * TC_UMMA_K), umma_desc_sw128(b_addr + k * TC_UMMA_K), TC_IDESC,
                            (kb | k) != 0);
                }
                umma_commit(empty_bar(s));                       // frees the stage when these MMAs have read it
                if (kb == k_blocks - 1) umma_commit(tmem_full_bar);   // accumulator complete
            }
        }
    } else {
        // ---- epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31; one thread = one ciphertext row -----------------
        const int quad = warp & 3;
        const int row = m_tile * TC_BM + quad * 32 + lane;
        mbar_wait(tmem_full_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int width = n_out + 1;
        constexpr int CPL = 32 / PL;                 // coefficients per 32-column load
#pragma unroll 1
        for (int c = 0; c < TC_BN / 32; c++) {
            uint32_t v[32];
            tmem_ld32x32(tmem_acc + ((uint32_t)(quad * 32) << 16) + (uint32_t)(c * 32), v);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (row < batch) {
#pragma unroll
                for (int e = 0; e < CPL; e++) {
                    const int j = n_tile * COEF + c * CPL + e;
                    if (j < width) {
                        T sum = 0;
#pragma unroll
                        for (int p = 0; p < PL; p++) sum += (T)(typename TorusTraits<T>::Signed)(int32_t)v[e * PL + p] << (8 * p);
                        T base = (j == body_col) ? in[(size_t)row * in_stride + n_in] : (T)0;
                        out[(size_t)row * out_stride + j] = base - sum;
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "n"(TC_BN));
#endif
}

// digits[ct][i*level + t] = t-th digit the decomposition emits for input coefficient i (levels l_ks -> 1, the key's row order)
template <typename T>
__global__ void ks_digits_kernel(const T *__restrict__ in, size_t in_stride, int8_t *__restrict__ digits, size_t kp, int n_in,
                                 int level, int base_log, int batch, int sub, int sub_bits) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (size_t)batch * n_in) return;
    const size_t ct = e / n_in;
    const int i = (int)(e % n_in);
    T state = decomp_init_state<T>(in[ct * in_stride + i], base_log, level);
    int8_t *dst = digits + ct * kp + (size_t)i * level * sub;
    if (sub == 1) {
        for (int t = 0; t < level; t++) dst[t] = (int8_t)decomp_next_digit<T>(state, base_log);
        return;
    }
    // wide digits (|D| <= 2^(base_log-1), base_log <= 7*sub): D = sum_s d_s * 2^(sub_bits*(sub-1-s)) with the low sub-digits
    // balanced in [-2^(sub_bits-1), 2^(sub_bits-1)) and the top one taking the rest (|d_0| <= 2^(base_log-1-sub_bits*(sub-1)) + 1)
    const int half = 1 << (sub_bits - 1), mask = (1 << sub_bits) - 1;
    for (int t = 0; t < level; t++) {
        int D = (int)(typename TorusTraits<T>::Signed)decomp_next_digit<T>(state, base_log);
        for (int s = sub - 1; s >= 1; s--) {
            const int lo = ((D + half) & mask) - half;
            dst[t * sub + s] = (int8_t)lo;
            D = (D - lo) >> sub_bits;
        }
        dst[t * sub] = (int8_t)D;
    }
}

template <typename T>
__global__ void ks_expand_rows_kernel(const T *__restrict__ src, T *__restrict__ dst, size_t n_in, int level, int sub, int sub_bits, int rev,
                                      size_t width) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t rows = n_in * (size_t)level * sub;
    if (e >= rows * width) return;
    const size_t row = e / width, j = e % width;
    const size_t i = row / ((size_t)level * sub);
    const int ts = (int)(row % ((size_t)level * sub)), t = ts / sub, s = ts % sub;
    const T v = src[(i * level + (rev ? level - 1 - t : t)) * width + j];
    dst[e] = (T)(v << (sub_bits * (sub - 1 - s)));
}

// planes[(j*PL + p)][r] = byte p of KSK[r][j]
template <typename T>
__global__ void ks_byteplanes_kernel(const T *__restrict__ ksk, uint8_t *__restrict__ planes, size_t rows, size_t kp, int width) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * (size_t)width) return;
    const size_t r = e / width;
    const int j = (int)(e % width);
    const T v = ksk[e];
    for (int p = 0; p < (int)sizeof(T); p++) planes[((size_t)j * sizeof(T) + p) * kp + r] = (uint8_t)(v >> (8 * p));
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// 2-D byte matrix [rows][kp] (kp contiguous), box = 128 bytes x box_rows, 128B swizzle, out-of-bounds reads give zeros
int make_map(CUtensorMap *map, const void *base, size_t rows, size_t k_extent, size_t kp, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return -2;
    cuuint64_t dims[2] = {(cuuint64_t)k_extent, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)kp};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -3;
}

}  // namespace

static_assert(sizeof(CUtensorMap) == sizeof(((KSTensorKey *)nullptr)->map_b), "tensor map storage size");

size_t ks_tc_padded_k(int n_in, int level) { return (((size_t)n_in * level) + 15) & ~(size_t)15; }

int ks_tc_supported(int bits, int n_in, int n_out, int level, int base_log) {
    (void)n_out;
    if (bits != 64 && bits != 32) return 0;
    if (base_log < 1 || base_log > 7 || level < 1 || level * base_log >= bits) return 0;
    // exactness of the int32 accumulation: rows * max|digit| * 255 < 2^31
    const double bound = (double)n_in * level * (double)(1 << (base_log - 1)) * 255.0;
    return bound < 2147483647.0 && encode_fn() != nullptr;
}

size_t ks_tc_key_bytes(int bits, int n_in, int n_out, int level) { return (size_t)(n_out + 1) * (bits / 8) * ks_tc_padded_k(n_in, level); }

int ks_tc_prepare_key(KSTensorKey *tk, const void *ksk, void *planes, int bits, int n_in, int n_out, int level, cudaStream_t st) {
    const size_t rows = (size_t)n_in * level, kp = ks_tc_padded_k(n_in, level);
    const int width = n_out + 1;
    cudaError_t e = cudaMemsetAsync(planes, 0, ks_tc_key_bytes(bits, n_in, n_out, level), st);
    if (e != cudaSuccess) return (int)e;
    const size_t total = rows * width;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64) ks_byteplanes_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)ksk, (uint8_t *)planes, rows, kp, width);
    else ks_byteplanes_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)ksk, (uint8_t *)planes, rows, kp, width);
    if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    tk->d_planes = planes;
    tk->kp = kp;
    return make_map(reinterpret_cast<CUtensorMap *>(tk->map_b), planes, (size_t)width * (bits / 8), rows, kp, TC_BN);
}

int ks_expand_rows_launch(const void *src, void *expanded, size_t n_in, int level, int sub, int sub_bits, int rev, size_t width, int bits,
                          cudaStream_t st) {
    const size_t total = n_in * (size_t)level * sub * width;
    if (total == 0) return 0;
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (bits == 64) ks_expand_rows_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)src, (uint64_t *)expanded, n_in, level, sub, sub_bits, rev, width);
    else ks_expand_rows_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)src, (uint32_t *)expanded, n_in, level, sub, sub_bits, rev, width);
    return (int)cudaGetLastError();
}

int ks_tc_launch(const KSLaunch &a, const KSTensorKey &tk, void *digits, cudaStream_t st) {
    if (a.batch <= 0) return 0;
    const size_t rows = (size_t)a.n_in * a.level * a.sub, kp = tk.kp;
    const int body_col = a.no_body ? a.n_out + 1 : (a.body_col < 0 ? a.n_out : a.body_col);   // n_out + 1: no column matches
    alignas(64) CUtensorMap map_a;
    int rc = make_map(&map_a, digits, (size_t)a.batch, rows, kp, TC_BM);
    if (rc) return rc;
    const size_t total = (size_t)a.batch * a.n_in;
    const unsigned dgrid = (unsigned)((total + 255) / 256);
    const int k_blocks = (int)((rows + TC_BK - 1) / TC_BK);
    const int pl = a.bits / 8;
    dim3 grid((unsigned)(((size_t)(a.n_out + 1) * pl + TC_BN - 1) / TC_BN), (unsigned)((a.batch + TC_BM - 1) / TC_BM));
    const CUtensorMap &map_b = *reinterpret_cast<const CUtensorMap *>(tk.map_b);
    cudaError_t e;
    if (a.bits == 64) {
        if ((e = ensure_dynamic_smem((const void *)ks_tc_kernel<uint64_t>, TC_SMEM)) != cudaSuccess) return (int)e;
        ks_digits_kernel<uint64_t><<<dgrid, 256, 0, st>>>((const uint64_t *)a.in, a.in_stride, (int8_t *)digits, kp, a.n_in, a.level, a.base_log, a.batch, a.sub, a.sub_bits);
        ks_tc_kernel<uint64_t><<<grid, TC_THREADS, TC_SMEM, st>>>(map_a, map_b, (const uint64_t *)a.in, a.in_stride, (uint64_t *)a.out,
                                                                 a.out_stride, a.n_in, a.n_out, k_blocks, a.batch, body_col);
    } else {
        if ((e = ensure_dynamic_smem((const void *)ks_tc_kernel<uint32_t>, TC_SMEM)) != cudaSuccess) return (int)e;
        ks_digits_kernel<uint32_t><<<dgrid, 256, 0, st>>>((const uint32_t *)a.in, a.in_stride, (int8_t *)digits, kp, a.n_in, a.level, a.base_log, a.batch, a.sub, a.sub_bits);
        ks_tc_kernel<uint32_t><<<grid, TC_THREADS, TC_SMEM, st>>>(map_a, map_b, (const uint32_t *)a.in, a.in_stride, (uint32_t *)a.out,
                                                                 a.out_stride, a.n_in, a.n_out, k_blocks, a.batch, body_col);
    }
    return (int)cudaGetLastError();
}

}  // namespace b200tfhe

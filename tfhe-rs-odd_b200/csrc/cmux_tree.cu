// cmux_tree.cu -- one layer of the CMUX tree of the WoP-PBS's vertical packing, batched.
//
// Reference: cmux_tree_memory_optimized, tfhe/src/core_crypto/fft_impl/fft64/crypto/wop_pbs/mod.rs:468-592; cmux,
// fft64/crypto/ggsw.rs:766-777: node = t0 + GGSW (x) (t1 - t0), an external product between two GLWE ciphertexts and a
// circuit-bootstrapped GGSW of the SAME input ciphertext.  The reference walks the tree of one look-up table depth-first
// with O(depth) buffers; here a layer of the trees of every (ciphertext, table) pair is one launch: CTA = one CMUX =
// the stages of one iteration of blind_rotate.cu (decompose, forward FFT, Fourier MAC, inverse FFT + accumulate) with the
// difference t1 - t0 in place of the rotate-and-subtract.
#include "br_core.cuh"
#include "kernels.h"

namespace b200tfhe {

template <typename T, int R, int K, int L, int BL> struct CmuxConfig {
    typedef BRCore<T, R, K, L, BL> C;
    static constexpr int K1 = K + 1;
    static constexpr int THREADS = K1 * R;
    static_assert(THREADS % 32 == 0, "teams must fill whole warps");
    static constexpr size_t FBUF_BYTES = (size_t)K1 * L * C::REGION * sizeof(cplx);
    static constexpr size_t POLY_BYTES = (size_t)K1 * C::N * sizeof(T);
    static constexpr size_t TW_BYTES = sizeof(TwTables<R>);
    static constexpr size_t SMEM = FBUF_BYTES + 2 * POLY_BYTES + TW_BYTES;
};

// grid = batch * n_luts * pairs.  CTA e = (ct, o, i):
//   t0 = in + (ct * ct_in_stride + (o * pairs + i) * 2) * glwe,  t1 = t0 + glwe      (ct_in_stride = 0: layer 0, the tables)
//   out + ((ct * n_luts + o) * pairs + i) * glwe  <-  t0 + GGSW[ct] (x) (t1 - t0),   GGSW[ct] = ggsw + ct * key_stride
template <typename T, int R, int K, int L, int BL>
__global__ void __launch_bounds__((K + 1) * R)
cmux_layer_kernel(const T *__restrict__ in, size_t ct_in_stride, T *__restrict__ out, const cplx *__restrict__ ggsw, size_t key_stride,
                  int n_luts, int pairs, const cplx *__restrict__ tw_fwd_, const cplx *__restrict__ tw_inv_) {
    typedef CmuxConfig<T, R, K, L, BL> CFG;
    typedef typename CFG::C C;
    constexpr int K1 = K + 1, N = C::N, REGION = C::REGION;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int r = threadIdx.x / R, lane = threadIdx.x % R;
    const size_t e = blockIdx.x;
    const int i = (int)(e % pairs);
    const size_t co = e / pairs, ct = co / n_luts;
    const int o = (int)(co % n_luts);
    const size_t glwe = (size_t)K1 * N;

    cplx *fbuf_ct = reinterpret_cast<cplx *>(smem_raw);
    cplx *fbuf_r = fbuf_ct + (size_t)r * L * REGION;
    T *acc_r = reinterpret_cast<T *>(smem_raw + CFG::FBUF_BYTES) + (size_t)r * N;
    T *diff_r = reinterpret_cast<T *>(smem_raw + CFG::FBUF_BYTES + CFG::POLY_BYTES) + (size_t)r * N;
    cplx *tw_s = reinterpret_cast<cplx *>(smem_raw + CFG::FBUF_BYTES + 2 * CFG::POLY_BYTES);
    constexpr int TW_HALF = TwTables<R>::ENTRIES * R;
    for (int q = threadIdx.x; q < TW_HALF; q += CFG::THREADS) {
        tw_s[q] = tw_fwd_[q];
        tw_s[TW_HALF + q] = tw_inv_[q];
    }
    const cplx(*twf)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s);
    const cplx(*twi)[R] = reinterpret_cast<const cplx(*)[R]>(tw_s + TW_HALF);

    const T *t0 = in + (ct * ct_in_stride + ((size_t)o * pairs + i) * 2) * glwe + (size_t)r * N;
    const T *t1 = t0 + glwe;
    for (int c = lane; c < N; c += R) {
        const T a = t0[c];
        acc_r[c] = a;
        diff_r[c] = t1[c] - a;                                   // wop_pbs/mod.rs:543-546
    }
    __syncthreads();

    const cplx *g = ggsw + ct * key_stride;
    plain_digits_fwd_pass1<T, R, K, L, BL>(lane, diff_r, fbuf_r, twf);
    __syncwarp();
    cplx F[R];
    C::template mac_prefetch_part<0>(lane, r, g, F);
#pragma unroll 1
    for (int t = 0; t < L; t++) {
        cplx x[R];
        C::fwd_pass2_load(lane, x, fbuf_r + t * REGION);
        __syncwarp();
        C::fwd_pass2_store(lane, x, fbuf_r + t * REGION);
    }
    if constexpr (R > C::CH) C::template mac_prefetch_part<1>(lane, r, g, F);
    __syncthreads();
    C::mac_finish(lane, r, g, fbuf_ct, F);
    __syncthreads();
    C::inv_pass1(lane, F, fbuf_r, twi);
    __syncwarp();
    C::inv_pass2_accumulate(lane, fbuf_r, acc_r);               // add_external_product_assign into the copy of t0 (:566-567)
    __syncwarp();
    T *dst = out + (co * pairs + i) * glwe + (size_t)r * N;
    for (int c = lane; c < N; c += R) dst[c] = acc_r[c];
}

// (T, R, K, L, BASE_LOG) of the circuit-bootstrapped GGSWs: WOPBS_PARAM_MESSAGE_2_CARRY_2_KS_PBS (cbs level 3, base 2^5)
//                                                        WOPBS_PARAM_MESSAGE_2_CARRY_1_KS_PBS / 4_CARRY_0 (k 3, N 512; cbs level 4, base 2^3)
//                                                        1_CARRY_2 / 3_CARRY_1 / 4_CARRY_1 / 6_CARRY_0 (cbs level 5, base 2^3)
#define CMUX_INSTANCES(X) X(uint64_t, 32, 1, 3, 5) X(uint64_t, 16, 3, 4, 3) X(uint64_t, 16, 3, 5, 3)

int cmux_layer_supported(const Params &pc) {
#define X(T, R, K, L, BL) \
    if (pc.bits == (int)sizeof(T) * 8 && pc.N == 2 * R * R && pc.k == K && pc.pbs_level == L && pc.pbs_base_log == BL) return R;
    CMUX_INSTANCES(X)
#undef X
    return 0;
}

int cmux_layer_launch(const Params &pc, const void *in, size_t ct_in_stride, void *out, const void *ggsw, size_t key_stride, int n_luts,
                      int pairs, size_t n_cmux, const void *tw_fwd, const void *tw_inv, cudaStream_t st) {
    if (n_cmux == 0) return 0;
#define X(T, R, K, L, BL)                                                                                                          \
    if (pc.bits == (int)sizeof(T) * 8 && pc.N == 2 * R * R && pc.k == K && pc.pbs_level == L && pc.pbs_base_log == BL) {              \
        typedef CmuxConfig<T, R, K, L, BL> CFG;                                                                                    \
        auto kern = cmux_layer_kernel<T, R, K, L, BL>;                                                                              \
        if (cudaError_t e = ensure_dynamic_smem((const void *)kern, CFG::SMEM)) return (int)e;                                      \
        kern<<<(unsigned)n_cmux, CFG::THREADS, CFG::SMEM, st>>>((const T *)in, ct_in_stride, (T *)out, (const cplx *)ggsw, key_stride, \
                                                                 n_luts, pairs, (const cplx *)tw_fwd, (const cplx *)tw_inv);         \
        return (int)cudaGetLastError();                                                                                            \
    }
    CMUX_INSTANCES(X)
#undef X
    return -1;
}

}  // namespace b200tfhe

// kernels.h -- internal launch interface between the C-ABI layer (api.cu) and the kernel translation units.
#pragma once
#include "common.cuh"
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <utility>

namespace b200tfhe {

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device, per-kernel setting: remember what was configured for the
// (kernel, current device) pair so that one process may drive several devices.  carveout_pct >= 0 also sets the preferred
// shared-memory carve-out.
inline cudaError_t ensure_dynamic_smem(const void *kern, size_t bytes, int carveout_pct = -1) {
    static std::mutex mu;
    static std::map<std::pair<const void *, int>, size_t> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> g(mu);
    size_t &have = done[std::make_pair(kern, dev)];
    if (bytes <= have) return cudaSuccess;
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess) return e;
    if (carveout_pct >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carveout_pct > 100 ? 100 : carveout_pct);
    have = bytes;
    return cudaSuccess;
}

enum { BR_MODE_LWE = 0, BR_MODE_GLWE = 1 };

struct BRLaunch {
    const void *lwe_in;    // [batch][lwe_stride] small-key LWE ciphertexts (device)
    size_t lwe_stride;     // in scalars, >= n+1
    const void *luts;      // [n_lut][(k+1)N] accumulators (device)
    const uint32_t *lut_idx;  // [batch] or nullptr (all use LUT 0)
    const void *bsk;       // device Fourier bootstrap key
    void *out;             // [batch][out_stride]
    size_t out_stride;
    int n, batch, mode;
    const void *tw_fwd, *tw_inv;
    // per-ciphertext keys (vertical packing over circuit-bootstrapped GGSWs): ciphertext ct reads its n GGSWs at
    // bsk + (ct / bsk_ct_div) * bsk_ct_stride (in complex values); 0 = one key shared by the batch
    size_t bsk_ct_stride = 0;
    int bsk_ct_div = 1;
    // kernel choice (b200tfhe_context_set_option "br_kernel" / "lat_max"): BR_AUTO picks by parameter set and batch size
    int force = 0;
    int lat_max = 2 * 148;
    int lat_cluster_max = 74;   // latency kernel: at most this many ciphertexts run on a two-CTA cluster each (two SMs per ciphertext)
    mutable int launches = 1;   // set by br_launch: kernels it actually launched (a tail wave may go to a second kernel)
    // multi-bit keys, batches of at most mb_lat_max ciphertexts: scratch for the per-ciphertext summed keys of the two-kernel
    // latency path (br_mb_lat_scratch_bytes); nullptr = always the fused kernel
    void *mb_scratch = nullptr;
    int mb_lat_max = 0;
};
enum { BR_AUTO = 0, BR_GENERIC = 1, BR_TMEM = 2, BR_LAT = 3, BR_GENERIC_1CT = 4, BR_WS = 5, BR_X16 = 6, BR_PAIR = 7 };

// blind_rotate.cu
int br_supported(const Params &p);  // returns team radix R (16/32) or 0
int br_launch(const Params &p, const BRLaunch &a, cudaStream_t st);
int bsk_convert_launch(const Params &p, const void *std_bsk, void *out, const void *tw_fwd, cudaStream_t st);

// blind_rotate_big.cu (N = 8192)
int br_big_supported(const Params &p);
int br_big_launch(const Params &p, const BRLaunch &a, cudaStream_t st);
int bsk_convert_big_launch(const Params &p, const void *std_bsk, void *out, const void *tables, cudaStream_t st);

// blind_rotate_tmem.cu (N = 2048, k = 1, l = 1: Fourier polynomials exchanged through tensor memory, 4 ct per CTA)
int br_tmem_supported(const Params &p);
int br_tmem_launch(const Params &p, const BRLaunch &a, cudaStream_t st);

// blind_rotate_x16.cu (boolean DEFAULT_PARAMETERS shape: half-warp teams, Fourier polynomials exchanged through tensor memory, 8 ct per CTA)
int br_x16_supported(const Params &p);
int br_x16_launch(const Params &p, const BRLaunch &a, cudaStream_t st);
void br_x16_set_variant(int v);

// blind_rotate_pair.cu (k = 3, N = 512, one level: six ciphertexts per CTA, two per warp)
int br_pair_supported(const Params &p);
int br_pair_launch(const Params &p, const BRLaunch &a, cudaStream_t st);

// blind_rotate_ws.cu (same parameter set; warp-specialised: Fourier warp + torus warp per ciphertext).  A measured-and-rejected
// experiment: compiled only with `make WS=1` (-DB200TFHE_BUILD_WS), never part of the default library
#ifdef B200TFHE_BUILD_WS
int br_ws_supported(const Params &p);
int br_ws_launch(const Params &p, const BRLaunch &a, cudaStream_t st);
#endif

// blind_rotate_lat.cu (same parameter set; ONE ciphertext per CTA spread over 128 threads: the single-PBS latency path)
int br_lat_supported(const Params &p);
int br_lat_launch(const Params &p, const BRLaunch &a, cudaStream_t st);
// multi-bit latency path: n_iter plain external products against per-ciphertext summed keys (see blind_rotate_mb.cu)
int br_lat_plain_supported(const Params &p);
int br_lat_plain_launch(const Params &p, const BRLaunch &a, int n_iter, size_t key_ct_stride, cudaStream_t st);

// blind_rotate_mb.cu (multi-bit bootstrap keys: 2^G GGSWs per group of G mask elements, n/G iterations)
int br_mb_supported(const Params &p, int grouping);   // returns team radix R or 0
int br_mb_launch(const Params &p, int grouping, const BRLaunch &a, cudaStream_t st);
// bytes of BRLaunch::mb_scratch the two-kernel latency path needs for `batch` ciphertexts; 0 = that path is not built for the set
size_t br_mb_lat_scratch_bytes(const Params &p, int grouping, int batch);
// key conversion of `npoly` standard-domain polynomials with an explicit team radix (blind_rotate.cu)
int bsk_convert_polys_launch(int R, int bits, size_t npoly, const void *std_bsk, void *out, const void *tw_fwd, cudaStream_t st);

// keyswitch.cu
struct KSLaunch {
    const void *in;   size_t in_stride;    // [batch][in_stride], n_in+1 used
    void *out;        size_t out_stride;   // [batch][out_stride], n_out+1 written
    const void *ksk;                       // device copy of the reference layout [n_in][l_ks][n_out+1]
    const void *colsum;                    // [n_out+1]  2^(b-1) * sum of all key rows (digit offset correction)
    int n_in, n_out, level, base_log, bits, batch;
    int body_col = -1;                     // output column that receives the input body; -1: the last one (n_out).  The packing
                                           // keyswitch (LWE -> GLWE) puts it on coefficient 0 of the body polynomial (k*N)
    // tensor-core path only (private functional packing keyswitch, base_log up to 7 * sub): every reference digit is split
    // EXACTLY into `sub` balanced sub-digits of sub_bits bits (most significant first) that meet key rows pre-shifted by
    // 2^(sub_bits * (sub-1-s)), so sum_s d_s * (K << ...) == digit * K mod 2^W and the int8 GEMM stays bit-exact
    int sub = 1, sub_bits = 0;
    bool no_body = false;                  // the input body is an ordinary decomposed row (n_in counts it); nothing is added
    mutable int launches = 1;              // set by ks_launch: kernels it launched (2 on the split-row path of tiny batches)
};
int ks_launch(const KSLaunch &a, cudaStream_t st);
int ks_colsum_launch(const void *ksk, void *colsum, int n_in, int n_out, int level, int base_log, int bits, cudaStream_t st);

// keyswitch_tc.cu (tcgen05 kind::i8 path: digits x byte-planes of the key)
struct KSTensorKey {
    void *d_planes = nullptr;              // [(n_out+1) * W/8][kp] bytes, kp = padded n_in * level
    size_t kp = 0;
    alignas(64) unsigned char map_b[128];  // CUtensorMap of d_planes (opaque here: kernels.h does not include cuda.h)
};
int ks_tc_supported(int bits, int n_in, int n_out, int level, int base_log);
size_t ks_tc_padded_k(int n_in, int level);
size_t ks_tc_key_bytes(int bits, int n_in, int n_out, int level);
int ks_tc_prepare_key(KSTensorKey *tk, const void *ksk, void *planes, int bits, int n_in, int n_out, int level, cudaStream_t st);
// expanded[i][t*sub + s][j] = src[i][rev ? level-1-t : t][j] << (sub_bits * (sub-1-s)): the key rows the sub-digit split meets
int ks_expand_rows_launch(const void *src, void *expanded, size_t n_in, int level, int sub, int sub_bits, int rev, size_t width, int bits,
                          cudaStream_t st);
int ks_tc_launch(const KSLaunch &a, const KSTensorKey &tk, void *digits, cudaStream_t st);   // digits: batch * kp bytes of scratch

// noalloc.cu: the same two launches with the digit scratch in a file-scope __device__ array (no cudaMalloc anywhere on the path)
constexpr size_t NOALLOC_KS_SCRATCH_BYTES = (size_t)4 << 20;   // == B200TFHE_NOALLOC_KS_SCRATCH_BYTES of the public header
int ks_tc_noalloc_launch(const KSLaunch &a, const KSTensorKey &tk, int device, cudaStream_t st, cudaError_t *cuda_err);

// wave_kernels.cu (device-resident waves over a ciphertext pool); WaveOp is layout-identical to b200tfhe_wave_op
struct WaveOp {
    uint32_t dst;
    uint32_t src[4];
    uint32_t src4;
    int64_t scale[4];
    uint64_t plaintext;
    int32_t lut;
    int32_t scale4;
};
int wave_gather_launch(const void *pool, size_t len, const WaveOp *d_ops, int n_ops, void *staging, uint32_t *d_lut_idx, int n_pbs,
                       int bits, cudaStream_t st);
int wave_scatter_launch(void *pool, size_t len, const WaveOp *d_ops, int n_ops, int n_pbs, const void *pbs_out, const void *staging,
                        int bits, cudaStream_t st);

// gadget_kernels.cu (multi-value bootstrap post-processing, accumulator packing: gadget/engine/bootstrapping.rs)
int glwe_mul_plain_extract_launch(const void *glwe, const void *polys, void *out, int k, int N, int n_polys, int batch, int bits,
                                  cudaStream_t st);
int pack_elements_launch(const void *in, void *out, int count, int p, size_t len, int batch, int bits, cudaStream_t st);
int pack_rotate_sum_launch(const void *G, void *out, int p, int k, int N, int batch, int bits, cudaStream_t st);

// wopbs_kernels.cu (element-wise steps of the bit extraction: fft64/crypto/wop_pbs/mod.rs:66-225)
int eb_update_shift_launch(void *buffer, const void *pbs, void *shifted, size_t len, int batch, int sh, uint64_t alpha, int bits,
                           cudaStream_t st);
int eb_store_launch(void *ks, void *out, size_t len, int batch, int nbits, int slot, int add_quarter, int bits, cudaStream_t st);
// circuit bootstrap glue (wop_pbs/mod.rs:243-444): bootstrap batch [batch][nbits (reversed)][levels] with body + q/4 and the
// accumulator index per entry; + alpha(level) on the bootstrap outputs; idx[e] = e % mod
int cbs_prep_launch(const void *bits_in, void *out, uint32_t *lut_idx, size_t len, int batch, int nbits, int levels, int bits, cudaStream_t st);
int cbs_post_launch(void *pbs, size_t len, size_t entries, int levels, int base_log, int bits, cudaStream_t st);
int mod_index_launch(uint32_t *idx, size_t total, uint32_t mod, cudaStream_t st);

// cmux_tree.cu (one layer of the vertical packing's CMUX tree: wop_pbs/mod.rs:468-592); pc carries (cbs_level, cbs_base_log)
int cmux_layer_supported(const Params &pc);
int cmux_layer_launch(const Params &pc, const void *in, size_t ct_in_stride, void *out, const void *ggsw, size_t key_stride, int n_luts,
                      int pairs, size_t n_cmux, const void *tw_fwd, const void *tw_inv, cudaStream_t st);

// external_product.cu (stand-alone add_external_product_assign / cmux_assign over a batch, every blind-rotation parameter set)
int external_product_supported(const Params &p);   // team radix (16/32/256) or 0
int external_product_launch(const Params &p, void *out, void *src, const void *ggsw_list, const uint32_t *idx, int cmux, int batch,
                            const void *tw_fwd, const void *tw_inv, cudaStream_t st);

// misc_kernels.cu
int modswitch_launch(const void *in, uint32_t *out, size_t count, int log2N, int bits, cudaStream_t st);
int sample_extract_launch(const void *glwe, void *lwe, int k, int N, int batch, int bits, cudaStream_t st);
int dfma_probe_launch(double *scratch, int grid, int iters, cudaStream_t st);
int gate_prologue_launch(int gate, const uint32_t *a, const uint32_t *b, uint32_t *out, int len, int batch, cudaStream_t st);

}  // namespace b200tfhe

/*
 * b200tfhe.h -- C ABI of the B200-native batched keyswitch + programmable-bootstrap library.
 *
 * This is the drop-in boundary for ONE path of TFHE-rs (reference: nicolasbon38/tfhe-rs-odd, a TFHE-rs
 * v0.5.0 fork): what a Rust `extern "C"` sys-crate would bind in place of the generic functions below.
 * Citations are relative to the reference's tfhe/src/.  Conventions mirror the reference's own C API
 * (c_api/utils.rs:3-27): every function returns int, 0 = success, non-zero = failure (the reference
 * turns panics into 1; here argument checks replace its assert!s); handles are opaque, allocated by the
 * library and freed by the matching *_destroy; out-handles are set to NULL before any work.
 *
 * Containers are the reference's flat layouts, passed as plain pointers:
 *   LWE ciphertext   [mask 0..dim-1, body]                         core_crypto/entities/lwe_ciphertext.rs:598-609
 *   LWE list         batch x (dim+1), contiguous                   entities/lwe_ciphertext_list.rs:10-17
 *   GLWE ciphertext  (k+1) polynomials of N coefficients, body last entities/glwe_ciphertext.rs:210-223
 *   LweBootstrapKey  [n][level 1..l][row 0..k][poly 0..k][N]        entities/lwe_bootstrap_key.rs:99-106,
 *                    (standard domain)                              entities/ggsw_ciphertext.rs:185-197
 *   LweKeyswitchKey  [n_in][l_ks (level l_ks first)][n_out+1]       entities/lwe_keyswitch_key.rs:77-108,
 *                                                                   algorithms/lwe_keyswitch_key_generation.rs:109-120
 * Only native power-of-two moduli (2^64: *_u64, 2^32: *_u32) are accepted, as the reference asserts
 * (fft_impl/fft64/crypto/bootstrap.rs:254, algorithms/lwe_keyswitch.rs:136-141).
 *
 * Pointer flavours: plain entry points take HOST pointers and stage host<->device copies themselves on the
 * context's stream (synchronous on return); *_dev entry points take DEVICE pointers, enqueue on the
 * context's stream and return without synchronising (call b200tfhe_synchronize).
 *
 * There is no CPU fallback: every compute entry point fails with a non-zero code if no CUDA device is usable.
 */
#ifndef B200TFHE_H
#define B200TFHE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200tfhe_context b200tfhe_context;
typedef struct b200tfhe_bsk b200tfhe_bsk; /* Fourier bootstrap key resident on the device: FourierLweBootstrapKey,
                                             fft_impl/fft64/crypto/bootstrap.rs:27-33 */
typedef struct b200tfhe_ksk b200tfhe_ksk; /* LweKeyswitchKey resident on the device */

/* parameter newtypes of core_crypto/commons/parameters.rs:40-236, flattened */
typedef struct {
    int32_t lwe_dimension;    /* n   */
    int32_t glwe_dimension;   /* k   */
    int32_t polynomial_size;  /* N   */
    int32_t pbs_base_log;
    int32_t pbs_level;
    int32_t ks_base_log;      /* 1 .. 28 (above 7: exact digit split on the tensor-core keyswitch) */
    int32_t ks_level;
    int32_t torus_bits;       /* 64 or 32 */
} b200tfhe_params;

/* boolean binary gates, boolean/engine/mod.rs:607-798 */
enum { B200TFHE_GATE_AND = 0, B200TFHE_GATE_NAND = 1, B200TFHE_GATE_OR = 2, B200TFHE_GATE_NOR = 3,
       B200TFHE_GATE_XOR = 4, B200TFHE_GATE_XNOR = 5 };

/* ---- context ------------------------------------------------------------------------------------------ */
/* One context = one CUDA device + one stream + a private workspace pool (the per-thread ShortintEngine /
 * ComputationBuffers of shortint/engine/mod.rs:23-69).  Calls on one context are serialised internally, but a context (and its
 * last-error text) belongs to ONE host thread at a time, like the reference's thread_local engines: use one context per thread;
 * key handles are immutable after upload and belong to the context that created them.
 * A host-pointer entry point that fails after it started copying drains the context's streams before it returns. */
int b200tfhe_context_create(int device, b200tfhe_context **out);
int b200tfhe_context_destroy(b200tfhe_context *ctx);
int b200tfhe_synchronize(b200tfhe_context *ctx);
/* last error text of this context (or of context creation when ctx == NULL); never NULL */
const char *b200tfhe_last_error(const b200tfhe_context *ctx);
/* the context's CUDA stream as a cudaStream_t cast to void* (for callers that interleave their own work) */
void *b200tfhe_context_stream(b200tfhe_context *ctx);
/* Tuning / test knobs of one context (they replace the environment switches of earlier builds; production needs none):
 *   "br_kernel"       0 auto (default) | 1 generic multi-ciphertext kernel | 2 tensor-memory kernel (N 2048, k 1, l 1) |
 *                     3 one-ciphertext-per-SM latency kernel | 4 generic kernel, one ciphertext per CTA | 5 warp-specialised
 *                     experiment (only in a `make WS=1` build).  A kernel that is not built for the key's parameter set makes
 *                     the blind rotation fail with an error -- the parity tests run every kernel on the same inputs this way.
 *                     6 = the half-warp-team tensor-memory kernel of the boolean shape (k 2, N 512, l 3).
 *   "x16_variant"     organisation of kernel 6 (process-wide): -1 default (= 12: all transforms first + TMA key ring), 4 = all transforms
 *                     first with the key from L2, 0 / 1 = level-by-level with the key from L2 / from a TMA-staged level block
 *   "lat_max"         largest batch that runs on the latency kernel in auto mode (default 296 = 2 x 148 SMs; 0 = never)
 *   "ks_tensor_core"  0: use the CUDA-core keyswitch even where the tensor-core key planes exist (default 1)
 *   "ks_tc_min"       smallest batch on the tensor-core keyswitch (default 1)
 *   "chunk_min"       minimum chunk (ciphertexts) of the H2D / compute / D2H pipeline of host-pointer calls (default 3552; 0 = off)
 * Unknown keys and out-of-range values are errors. */
int b200tfhe_context_set_option(b200tfhe_context *ctx, const char *key, int64_t value);
/* 1 if blind rotation is implemented for this parameter set on this build, else 0 */
int b200tfhe_params_supported(const b200tfhe_params *p);

/* ---- keys --------------------------------------------------------------------------------------------- */
/* Replaces par_convert_standard_lwe_bootstrap_key_to_fourier (algorithms/lwe_bootstrap_key_conversion.rs:99-152):
 * uploads a standard-domain LweBootstrapKey and runs the forward negacyclic FFT on the device into the
 * device's own Fourier layout.  uses p->{lwe_dimension, glwe_dimension, polynomial_size, pbs_*, torus_bits}. */
int b200tfhe_bsk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *std_bsk, b200tfhe_bsk **out);
int b200tfhe_bsk_upload_u32(b200tfhe_context *ctx, const b200tfhe_params *p, const uint32_t *std_bsk, b200tfhe_bsk **out);
/* same, standard-domain key already on the device */
int b200tfhe_bsk_convert_dev(b200tfhe_context *ctx, const b200tfhe_params *p, const void *d_std_bsk, b200tfhe_bsk **out);
/* allocate an uninitialised device Fourier key (to be filled by a peer copy / NCCL broadcast of another
 * GPU's key: the Fourier layout is identical on every B200) */
int b200tfhe_bsk_alloc(b200tfhe_context *ctx, const b200tfhe_params *p, b200tfhe_bsk **out);
/* device address and size in bytes of the Fourier key ( n*l*(k+1)^2*(N/2) complex f64 ) */
int b200tfhe_bsk_device_buffer(const b200tfhe_bsk *bsk, void **d_ptr, size_t *bytes);
int b200tfhe_bsk_destroy(b200tfhe_bsk *bsk);

/* ---- multi-bit bootstrap keys ----------------------------------------------------------------
 * LweMultiBitBootstrapKey (core_crypto/entities/lwe_multi_bit_bootstrap_key.rs; generation
 * algorithms/lwe_multi_bit_bootstrap_key_generation.rs:87-172): standard-domain layout
 * [n/G][2^G][level 1..l][row 0..k][poly 0..k][N], G = LweBskGroupingFactor.  Replaces
 * par_convert_standard_lwe_multi_bit_bootstrap_key_to_fourier (algorithms/lwe_multi_bit_bootstrap_key_conversion.rs).
 * The returned handle is an ordinary b200tfhe_bsk: every pbs / blind_rotate / ks_pbs / pbs_ks / wave entry point accepts
 * it and then computes multi_bit_(deterministic_)programmable_bootstrap_lwe_ciphertext
 * (algorithms/lwe_multi_bit_programmable_bootstrapping.rs:1035-1224) with n/G sequential iterations, as shortint
 * dispatches on ShortintBootstrappingKey::{Classic, MultiBit} (shortint/server_key/mod.rs:798-851).  The reference's
 * thread_count / deterministic_execution knobs have no counterpart: groups are always consumed in ascending order.
 * Only 64-bit sets exist in the reference (shortint/parameters/multi_bit.rs). */
int b200tfhe_multibit_bsk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping_factor,
                                     const uint64_t *std_bsk, b200tfhe_bsk **out);
int b200tfhe_multibit_bsk_convert_dev(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping_factor,
                                      const void *d_std_bsk, b200tfhe_bsk **out);
int b200tfhe_multibit_bsk_alloc(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping_factor, b200tfhe_bsk **out);
/* Upload a bootstrap key that is already in the Fourier domain, in the reference's serialised order
 * (FourierPolynomialList, fft_impl/fft64/math/fft/mod.rs:588-717: [ggsw][level][row][col] polynomials, each N/2 complex f64
 * in natural frequency order, re/im interleaved; values as forward_as_torus produces them).  grouping_factor 1 = classic
 * FourierLweBootstrapKey, >= 2 = FourierLweMultiBitBootstrapKey.  What deserialising a shortint ServerKey needs. */
int b200tfhe_bsk_upload_fourier(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping_factor, const double *fourier,
                                b200tfhe_bsk **out);
/* 1 for a classic key, G for a multi-bit key */
uint32_t b200tfhe_bsk_grouping_factor(const b200tfhe_bsk *bsk);
/* 1 if the multi-bit blind rotation is built for this parameter set and grouping factor, else 0 */
int b200tfhe_multibit_params_supported(const b200tfhe_params *p, uint32_t grouping_factor);

/* uploads an LweKeyswitchKey (input dim k*N, output dim n); uses p->{lwe_dimension, glwe_dimension,
 * polynomial_size, ks_*, torus_bits} */
int b200tfhe_ksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *ksk, b200tfhe_ksk **out);
int b200tfhe_ksk_upload_u32(b200tfhe_context *ctx, const b200tfhe_params *p, const uint32_t *ksk, b200tfhe_ksk **out);
int b200tfhe_ksk_alloc(b200tfhe_context *ctx, const b200tfhe_params *p, b200tfhe_ksk **out);
/* device address/size of the key rows; after filling it from a peer call b200tfhe_ksk_finalize */
int b200tfhe_ksk_device_buffer(const b200tfhe_ksk *ksk, void **d_ptr, size_t *bytes);
int b200tfhe_ksk_finalize(b200tfhe_context *ctx, b200tfhe_ksk *ksk);
int b200tfhe_ksk_destroy(b200tfhe_ksk *ksk);

/* ---- batched operations, host pointers ------------------------------------------------------------------ */
/* keyswitch_lwe_ciphertext (algorithms/lwe_keyswitch.rs:96-170) over a list:
 * in: batch x (kN+1), out: batch x (n+1).  Bit-exact. */
int b200tfhe_keyswitch_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const uint64_t *in, uint64_t *out, uint32_t batch);
int b200tfhe_keyswitch_batch_u32(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const uint32_t *in, uint32_t *out, uint32_t batch);

/* programmable_bootstrap_lwe_ciphertext(_mem_optimized) (algorithms/lwe_programmable_bootstrapping.rs:1017-1113)
 * over a list: in: batch x (n+1) under the small key, out: batch x (kN+1).
 * luts: n_luts accumulators, each a GLWE of (k+1)N scalars; lut_idx[b] selects ciphertext b's accumulator
 * (NULL = all use accumulator 0). */
int b200tfhe_pbs_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const uint64_t *in, uint64_t *out,
                           const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_pbs_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const uint32_t *in, uint32_t *out,
                           const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);

/* programmable_bootstrap_lwe_ciphertext_without_sample_extract_mem_optimized (fork entry,
 * lwe_programmable_bootstrapping.rs:1115-1162) / blind_rotate_assign (:232-289): out: batch x (k+1)N GLWE */
int b200tfhe_blind_rotate_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const uint64_t *in, uint64_t *out_glwe,
                                    const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_blind_rotate_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const uint32_t *in, uint32_t *out_glwe,
                                    const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);

/* shortint keyswitch_programmable_bootstrap_assign core (shortint/server_key/mod.rs:783-857):
 * KS then PBS; in/out: batch x (kN+1) */
int b200tfhe_ks_pbs_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                              uint64_t *out, const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_ks_pbs_batch_u32(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint32_t *in,
                              uint32_t *out, const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
/* programmable_bootstrap_keyswitch_assign core (:859-933) / boolean bootstrap_keyswitch
 * (boolean/engine/bootstrapping.rs:299-344): PBS then KS; in/out: batch x (n+1) */
int b200tfhe_pbs_ks_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                              uint64_t *out, const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_pbs_ks_batch_u32(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint32_t *in,
                              uint32_t *out, const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);

/* add_external_product_assign(_mem_optimized) (algorithms/lwe_programmable_bootstrapping.rs:309-475 ->
 * fft_impl/fft64/crypto/ggsw.rs:477-598) over a list:  out[b] <- out[b] + ggsw[ggsw_idx[b]] (x) glwe[b].
 * The FourierGgswCiphertext list is a bootstrap-key handle (a classic LweBootstrapKey IS a list of lwe_dimension GGSW
 * ciphertexts, entities/lwe_bootstrap_key.rs:99-106): upload stand-alone GGSWs with b200tfhe_bsk_upload_* and
 * p->lwe_dimension = their count.  ggsw_idx[b] < lwe_dimension selects ciphertext b's GGSW (NULL: all use GGSW 0).
 * out (read and written) and glwe: batch x (k+1)N.  This is one blind-rotation iteration without the rotate-and-subtract;
 * the result differs from the reference's f64 path only by FFT rounding (per-CMUX bound). */
int b200tfhe_external_product_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *ggsw, const uint32_t *ggsw_idx, uint64_t *out,
                                        const uint64_t *glwe, uint32_t batch);
int b200tfhe_external_product_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *ggsw, const uint32_t *ggsw_idx, uint32_t *out,
                                        const uint32_t *glwe, uint32_t batch);
/* cmux_assign(_mem_optimized) (lwe_programmable_bootstrapping.rs:552-765 -> ggsw.rs:766-777) over a list, both operands
 * updated exactly as the reference updates them:  ct1[b] <- ct1[b] - ct0[b];  ct0[b] <- ct0[b] + ggsw[ggsw_idx[b]] (x) ct1[b]
 * (ct0 then encrypts ct1's old message where the GGSW encrypts 1, its own where it encrypts 0). */
int b200tfhe_cmux_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *ggsw, const uint32_t *ggsw_idx, uint64_t *ct0, uint64_t *ct1,
                            uint32_t batch);
int b200tfhe_cmux_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *ggsw, const uint32_t *ggsw_idx, uint32_t *ct0, uint32_t *ct1,
                            uint32_t batch);

/* boolean binary gate over a list (boolean/engine/mod.rs:607-798 + Bootstrapper::apply_bootstrapping_pattern,
 * engine/bootstrapping.rs:392-401): linear combination, then PBS->KS (pbs_first != 0, DEFAULT_PARAMETERS) or
 * KS->PBS (pbs_first == 0, *_KS_PBS parameters) with the constant accumulator of bootstrapping.rs:59-60.
 * a, b, out: batch x (dim+1), dim = n when pbs_first else kN */
int b200tfhe_gate_batch_u32(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, int gate, int pbs_first,
                            const uint32_t *a, const uint32_t *b, uint32_t *out, uint32_t batch);

/* stand-alone bit-exact stages (fused inside the PBS; exposed for parity checks and custom chains) */
/* fast_pbs_modulus_switch, fft_impl/common.rs:26-43 (offset 0, lut_count_log 0): out[i] in [0, 2N] */
int b200tfhe_modulus_switch_u64(b200tfhe_context *ctx, const uint64_t *in, uint32_t *out, size_t count, uint32_t log2_poly_size);
int b200tfhe_modulus_switch_u32(b200tfhe_context *ctx, const uint32_t *in, uint32_t *out, size_t count, uint32_t log2_poly_size);
/* extract_lwe_sample_from_glwe_ciphertext with MonomialDegree(0), algorithms/glwe_sample_extraction.rs:91-147 */
int b200tfhe_sample_extract_batch_u64(b200tfhe_context *ctx, const uint64_t *glwe, uint64_t *lwe, uint32_t glwe_dimension,
                                      uint32_t polynomial_size, uint32_t batch);
int b200tfhe_sample_extract_batch_u32(b200tfhe_context *ctx, const uint32_t *glwe, uint32_t *lwe, uint32_t glwe_dimension,
                                      uint32_t polynomial_size, uint32_t batch);

/* ---- batched operations, device pointers (asynchronous on the context's stream) ---------------------------- */
int b200tfhe_keyswitch_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const void *d_in, void *d_out, uint32_t batch);
int b200tfhe_pbs_batch_dev(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const void *d_in, void *d_out, const void *d_luts,
                           const uint32_t *d_lut_idx, uint32_t batch);
int b200tfhe_blind_rotate_batch_dev(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const void *d_in, void *d_out_glwe,
                                    const void *d_luts, const uint32_t *d_lut_idx, uint32_t batch);
int b200tfhe_ks_pbs_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const void *d_in,
                              void *d_out, const void *d_luts, const uint32_t *d_lut_idx, uint32_t batch);
int b200tfhe_pbs_ks_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const void *d_in,
                              void *d_out, const void *d_luts, const uint32_t *d_lut_idx, uint32_t batch);

/* Allocation-free variant of b200tfhe_keyswitch_batch_dev.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and result bits as b200tfhe_keyswitch_batch_dev (tensor-core keyswitch), but the
 * int8 digit scratch is a fixed file-scope __device__ array of B200TFHE_NOALLOC_KS_SCRATCH_BYTES (held by every GPU that
 * loads the library until the process exits) instead of the context's growable workspace: the call performs no cudaMalloc /
 * cudaFree.  batch * padded(k*N*ks_level*sub) must fit the array (409 ciphertexts at PARAM_MESSAGE_2_CARRY_2), larger
 * batches are rejected with an error; the key must have tensor-core byte planes (every parameter set of the reference has).
 * Every call on a device shares the one array, which keeps the previous call's contents (nothing in it has to start at zero:
 * each call rewrites all it reads): calls that OVERLAP on the GPU -- two contexts / streams on one device -- race on it. */
#define B200TFHE_NOALLOC_KS_SCRATCH_BYTES (4u << 20)
int b200tfhe_keyswitch_batch_noalloc_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const void *d_in, void *d_out, uint32_t batch);

/* ---- gadget extras of the fork (tfhe/src/gadget/engine/bootstrapping.rs), 64-bit torus ---------------------
 * LwePackingKeyswitchKey (core_crypto/entities/lwe_packing_keyswitch_key.rs; generation
 * algorithms/lwe_packing_keyswitch_key_generation.rs:60-150): [kN][l_ks (level l_ks first)][(k+1)N], input key = the big LWE
 * key, output key = the GLWE key, decomposition = p->ks_* (as the gadget server key builds it, bootstrapping.rs:355-363). */
typedef struct b200tfhe_pksk b200tfhe_pksk;
typedef struct b200tfhe_pfpksk b200tfhe_pfpksk;   /* circuit-bootstrap private functional packing keyswitch key list */
int b200tfhe_pksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *pksk, b200tfhe_pksk **out);
int b200tfhe_pksk_destroy(b200tfhe_pksk *pksk);
/* keyswitch_lwe_ciphertext_into_glwe_ciphertext (algorithms/lwe_packing_keyswitch.rs:100-187) over a list:
 * in: batch x (kN+1), out: batch x (k+1)N.  Bit-exact.  Runs on the tensor-core keyswitch (same integer GEMM). */
int b200tfhe_packing_keyswitch_batch_u64(b200tfhe_context *ctx, const b200tfhe_pksk *pksk, const uint64_t *in, uint64_t *out_glwe,
                                         uint32_t batch);
/* out[b][i] = extract_lwe_sample_from_glwe_ciphertext( glwe[b] * polys[i], 0 ): the post-processing of mvb_bootstrap
 * (bootstrapping.rs:596-611: polynomial_karatsuba_wrapping_mul by the plaintext polynomial v_i, then sample extraction).
 * glwe: batch x (k+1)N, polys: n_polys x N, out: batch x n_polys x (kN+1).  Bit-exact. */
int b200tfhe_glwe_mul_plain_extract_batch_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *glwe, const uint64_t *polys,
                                              uint32_t n_polys, uint64_t *out, uint32_t batch);
/* mvb_bootstrap (bootstrapping.rs:567-622) over a list: [keyswitch when ksk != NULL (Big key choice: in is batch x (kN+1),
 * else batch x (n+1))] -> blind rotation of the common-factor accumulator (bootstrap_common_factor :451-500) ->
 * n_polys polynomial products + sample extractions.  out: batch x n_polys x (kN+1). */
int b200tfhe_mvb_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                           const uint64_t *common_acc, const uint64_t *polys, uint32_t n_polys, uint64_t *out, uint32_t batch);
/* pack_into_new_accumulator (bootstrapping.rs:690-773) over a list of accumulators: lwes: batch x count x (kN+1) with
 * 1 <= count <= p_mod (missing ones are zero ciphertexts), p_mod odd; out_glwe: batch x (k+1)N.  Bit-exact. */
int b200tfhe_pack_into_accumulator_batch_u64(b200tfhe_context *ctx, const b200tfhe_pksk *pksk, const uint64_t *lwes, uint32_t count,
                                             uint32_t p_mod, uint64_t *out_glwe, uint32_t batch);

/* ---- bit extraction, the first half of the WoP-PBS -------------------------------------------
 * Replaces extract_bits_from_lwe_ciphertext_mem_optimized (core_crypto/algorithms/lwe_wopbs.rs:321-364 ->
 * fft64/crypto/wop_pbs/mod.rs:66-225) over a ciphertext list.  in: batch x (kN+1) under the big key; out: batch x nbits x (n+1)
 * under the small key, the most significant extracted bit FIRST, every bit scaled by q/2.  Bits delta_log ..
 * delta_log + nbits - 1 of the plaintext are extracted from the lowest upwards: nbits keyswitches and nbits - 1 bootstraps
 * per ciphertext, all rounds device-resident.  delta_log + nbits <= 64 (the reference's debug_assert :86-94).  The first
 * extracted ciphertext (index nbits - 1) is bit-identical to the reference's; the later ones depend on bootstraps. */
int b200tfhe_extract_bits_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                                    uint32_t delta_log, uint32_t nbits, uint64_t *out, uint32_t batch);

/* ---- circuit bootstrap + vertical packing, the second half of the WoP-PBS ---------------------
 * Key list of generate_circuit_bootstrap_lwe_pfpksk_list (core_crypto/algorithms/lwe_wopbs.rs:80-160) in the reference layout
 * LwePrivateFunctionalPackingKeyswitchKeyList: [k+1 keys][kN+1 input elements, the body's last][level 1..pfks_level][(k+1)N].
 * p gives k, N and the torus width; pfks_base_log up to 28. */
int b200tfhe_cbs_pfpksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t pfks_base_log, uint32_t pfks_level,
                                   const uint64_t *list, b200tfhe_pfpksk **out);
int b200tfhe_cbs_pfpksk_destroy(b200tfhe_pfpksk *k);
/* private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext
 * (core_crypto/algorithms/lwe_private_functional_packing_keyswitch.rs:8-89) with key `key_index` of the list over a batch:
 * in batch x (kN+1), out_glwe batch x (k+1)N.  Bit-exact (tcgen05 kind::i8 GEMM over exactly split digits). */
int b200tfhe_pfpks_batch_u64(b200tfhe_context *ctx, const b200tfhe_pfpksk *pf, uint32_t key_index, const uint64_t *in, uint64_t *out_glwe,
                             uint32_t batch);
/* circuit_bootstrap_boolean_vertical_packing_lwe_ciphertext_list_mem_optimized (lwe_wopbs.rs:651-704 ->
 * fft64/crypto/wop_pbs/mod.rs:647-853) over a batch.  bits_in: batch x nbits x (n+1), the output of extract_bits (most
 * significant bit first, each scaled by q/2); luts: n_luts x polys_per_lut x N (big_lut_as_polynomial_list; polys_per_lut = 2^t,
 * t <= nbits, nbits - t <= log2 N); out: batch x n_luts x (kN+1) under the big key, out[b][o] encrypts coefficient `index` of
 * table o, index = the nbits bits of b read as a number (polynomial index = its t most significant bits).
 * Per input: nbits * cbs_level bootstraps (homomorphic_shift_boolean :369-444), (k+1) private functional packing
 * keyswitches each, Fourier conversion of the GGSWs; per (input, table): the CMUX tree over the table's polynomials
 * (cmux_tree_memory_optimized :468-592, one launch per layer) and a blind rotation over the input's own remaining GGSWs
 * (:866-892) with sample extraction.  Batches above 1024 inputs are processed in slices (bounded device workspace). */
int b200tfhe_cbs_vertical_packing_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const b200tfhe_pfpksk *pf, const uint64_t *bits_in,
                                            uint32_t nbits, const uint64_t *luts, uint32_t n_luts, uint32_t polys_per_lut,
                                            uint32_t cbs_base_log, uint32_t cbs_level, uint64_t *out, uint32_t batch);

/* ---- device-resident waves over a ciphertext pool ---------------------------------------------
 * The reference evaluates a shortint / integer / Trivium circuit as rayon-parallel layers of
 *   leveled prologue (unchecked_add, unchecked_scalar_mul: core_crypto/algorithms/lwe_linear_algebra.rs:191-215,556-572;
 *   bivariate packing lhs*factor + rhs: shortint/server_key/bivariate_pbs.rs:167-182)  +  apply_lookup_table
 * (integer/server_key/radix_parallel/mul.rs:329-424, apps/trivium/src/trivium/trivium_shortint.rs:81-168).  A pool keeps
 * the ciphertexts of such a circuit in HBM; b200tfhe_wave_run executes one layer:
 *     pool[dst] <- LUT_lut( sum_k scale[k] * pool[src[k]] + (0,...,0,plaintext) )      for every op (up to five terms), lut < 0: no bootstrap
 * Every op of a wave reads the pool as it was before the wave.  The first n_pbs ops are the bootstrapped ones (lut >= 0),
 * the rest are purely linear (lut < 0).  Asynchronous on the context's stream. */
typedef struct b200tfhe_pool b200tfhe_pool;
typedef struct {
    uint32_t dst;          /* pool slot written */
    uint32_t src[4];       /* pool slots read (ignored where scale == 0) */
    uint32_t src4;         /* fifth term (a column sum of radix blocks takes (p - 1)/(m - 1) = 5 terms at MESSAGE_2_CARRY_2:
                            * integer/server_key/radix_parallel/sum.rs); ignored where scale4 == 0 */
    int64_t scale[4];      /* cleartext multipliers, wrapping */
    uint64_t plaintext;    /* added to the body (already encoded, e.g. value * delta) */
    int32_t lut;           /* accumulator index into the pool's LUT table, or < 0 */
    int32_t scale4;        /* multiplier of the fifth term; 0 = no fifth term (these two words were reserved-zero before) */
} b200tfhe_wave_op;
/* big != 0: slots hold kN+1 scalars and a bootstrap is keyswitch -> PBS (shortint *_KS_PBS sets); big == 0: n+1 scalars,
 * PBS -> keyswitch (boolean DEFAULT_PARAMETERS) */
int b200tfhe_pool_create(b200tfhe_context *ctx, const b200tfhe_params *p, int big, uint32_t capacity, b200tfhe_pool **out);
int b200tfhe_pool_destroy(b200tfhe_pool *pool);
int b200tfhe_pool_write(b200tfhe_context *ctx, b200tfhe_pool *pool, uint32_t first, uint32_t count, const void *host_cts);
int b200tfhe_pool_read(b200tfhe_context *ctx, const b200tfhe_pool *pool, uint32_t first, uint32_t count, void *host_cts);
int b200tfhe_pool_device_buffer(const b200tfhe_pool *pool, void **d_ptr, size_t *bytes);
/* accumulators (GLWE, (k+1)N scalars each) used by the pool's waves; copied to the device */
int b200tfhe_pool_set_luts(b200tfhe_context *ctx, b200tfhe_pool *pool, const void *host_luts, uint32_t n_luts);
int b200tfhe_wave_run(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, b200tfhe_pool *pool,
                      const b200tfhe_wave_op *ops, uint32_t n_ops, uint32_t n_pbs);

/* A pre-scheduled circuit in one call: n_waves waves, concatenated in `ops` (wave w: n_ops[w] operations, the first n_pbs[w]
 * bootstrapped), executed back to back on the context's stream with the semantics of b200tfhe_wave_run per wave.  The host
 * scheduler (rayon layers in the reference: integer/server_key/radix_parallel/mul.rs:355-404) runs once, the plan is replayed. */
int b200tfhe_wave_sequence_run(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, b200tfhe_pool *pool,
                               const b200tfhe_wave_op *ops, const uint32_t *n_ops, const uint32_t *n_pbs, uint32_t n_waves);

/* ---- several GPUs of one node behind one handle (SURVEY 8(b) ctx_create(n_gpus), 8(e)) ------------------------------------------
 * The reference parallelises this path with rayon over independent ciphertexts (tfhe/benches/core_crypto/pbs_bench.rs:517-532,
 * integer/server_key/radix_parallel/mul.rs:355-404); here ONE host batch is split contiguously over the GPUs (remainder to the low
 * devices), one host thread + stream + workspace per GPU, results written back to the caller's buffer at the same indices, no
 * per-batch collective.  Keys are uploaded and converted once on device 0 and fanned out device-to-device (cudaMemcpyPeerAsync over
 * NVLink / NVSwitch).  n_gpus <= 0: every visible device.  Either key pointer may be NULL (keyswitch-only / bootstrap-only sets).
 * Host pointers throughout (pin them for full PCIe bandwidth); calls are synchronous. */
typedef struct b200tfhe_multi_context b200tfhe_multi_context;
typedef struct b200tfhe_multi_keys b200tfhe_multi_keys;
int b200tfhe_multi_context_create(int n_gpus, b200tfhe_multi_context **out);
int b200tfhe_multi_context_destroy(b200tfhe_multi_context *m);
int b200tfhe_multi_context_gpus(const b200tfhe_multi_context *m);
/* the single-GPU context of device `gpu` (owned by the multi context), or NULL */
b200tfhe_context *b200tfhe_multi_context_get(b200tfhe_multi_context *m, int gpu);
const char *b200tfhe_multi_last_error(const b200tfhe_multi_context *m);
int b200tfhe_multi_keys_upload_u64(b200tfhe_multi_context *m, const b200tfhe_params *p, const uint64_t *std_bsk, const uint64_t *ksk,
                                   b200tfhe_multi_keys **out);
int b200tfhe_multi_keys_upload_u32(b200tfhe_multi_context *m, const b200tfhe_params *p, const uint32_t *std_bsk, const uint32_t *ksk,
                                   b200tfhe_multi_keys **out);
int b200tfhe_multi_keys_destroy(b200tfhe_multi_keys *k);
/* the batched operations of the single-GPU section, same buffers and semantics, sharded over the GPUs */
int b200tfhe_multi_keyswitch_batch_u64(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint64_t *in, uint64_t *out, uint32_t batch);
int b200tfhe_multi_keyswitch_batch_u32(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint32_t *in, uint32_t *out, uint32_t batch);
int b200tfhe_multi_pbs_batch_u64(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint64_t *in, uint64_t *out,
                                 const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_multi_pbs_batch_u32(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint32_t *in, uint32_t *out,
                                 const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_multi_ks_pbs_batch_u64(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint64_t *in, uint64_t *out,
                                    const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_multi_ks_pbs_batch_u32(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint32_t *in, uint32_t *out,
                                    const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_multi_pbs_ks_batch_u64(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint64_t *in, uint64_t *out,
                                    const uint64_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);
int b200tfhe_multi_pbs_ks_batch_u32(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const uint32_t *in, uint32_t *out,
                                    const uint32_t *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch);

/* sustained FP64 FMA rate of the device in TFLOP/s (register-only probe kernel, best of repeated runs over
 * about duration_ms): the measured denominator of the blind rotation's FP64 roofline */
int b200tfhe_probe_fp64_tflops(b200tfhe_context *ctx, double duration_ms, double *tflops);

/* number of kernel launches issued through this context so far (bench.py's gpu_launches) */
uint64_t b200tfhe_launch_count(const b200tfhe_context *ctx);

#ifdef __cplusplus
}
#endif
#endif

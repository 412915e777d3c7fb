/*
 * tfhe_oracle.h -- CPU restatement of the reference's KS+PBS path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.  The CUDA library never links or calls it.
 *
 * PARITY STATUS: the integer stages (decomposer, modulus switch, monomial ops,
 * sample extract, keyswitch) are pinned by the reference's own known-answer
 * values (see tests/test_oracle_kat.py).  The raw f64 blind-rotation phase is
 * "parity unpinned": the reference holds no ciphertext-level golden vectors and
 * its FFT lives in the un-vendored crate concrete-fft 0.3.0 (tfhe/Cargo.toml:59);
 * it is pinned only through the reference's own tolerance tests
 * (fft_impl/fft64/math/fft/tests.rs:9-222) and decrypt-and-compare tests.
 *
 * All scalars are carried in uint64_t; `bits` (32 or 64) selects the torus width
 * and every result is reduced mod 2^bits (native power-of-two moduli only, as
 * the reference asserts in lwe_keyswitch.rs:136-141 and bootstrap.rs:254).
 *
 * Citations are relative to /root/reference/tfhe/src/.
 */
#ifndef TFHE_ORACLE_H
#define TFHE_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int n;            /* small LWE dimension */
    int k;            /* GLWE dimension */
    int N;            /* polynomial size */
    int pbs_base_log, pbs_level;
    int ks_base_log, ks_level;
    int bits;         /* 32 or 64 */
    double lwe_std, glwe_std;
} orc_params;

/* ---- integer primitives (bit-exact contracts) ---- */
/* commons/math/decomposition/decomposer.rs:99-119 */
uint64_t orc_closest_representable(uint64_t x, int base_log, int level, int bits);
/* commons/math/decomposition/iter.rs:37-50,120-127: out[0] = level `level` ... out[level-1] = level 1 */
void orc_decompose(uint64_t x, int base_log, int level, int bits, uint64_t *out_digits);
/* fft_impl/common.rs:26-43 (offset 0, lut_count_log 0) */
uint64_t orc_modswitch(uint64_t x, int log2N, int bits);
/* algorithms/polynomial_algorithms.rs:315-354 */
void orc_monomial_div(uint64_t *out, const uint64_t *in, int N, uint64_t deg, int bits);
/* algorithms/polynomial_algorithms.rs:425-490 : out = in*X^deg - in */
void orc_monomial_mul_sub(uint64_t *out, const uint64_t *in, int N, uint64_t deg, int bits);
/* algorithms/glwe_sample_extraction.rs:91-147, nth = 0 */
void orc_sample_extract0(uint64_t *lwe_out, const uint64_t *glwe, int k, int N, int bits);
/* algorithms/lwe_keyswitch.rs:96-170 */
void orc_keyswitch(const orc_params *p, const uint64_t *ksk, const uint64_t *in, uint64_t *out);

/* ---- negacyclic f64 FFT wrapper (fft_impl/fft64/math/fft/mod.rs) ---- */
/* forward_as_integer :220-239,496-515 ; out is M=N/2 complex (re,im interleaved), bit-reversed order */
void orc_fft_forward_integer(int N, const uint64_t *poly, int bits, double *out);
/* forward_as_torus :197-218 */
void orc_fft_forward_torus(int N, const uint64_t *poly, int bits, double *out);
/* add_backward_as_torus :285-326,539-557 ; fourier is destroyed */
void orc_fft_add_backward_torus(int N, uint64_t *poly, int bits, double *fourier);
/* schoolbook negacyclic product, reference test helper fft/tests.rs:83-100 */
void orc_negacyclic_schoolbook(int N, const uint64_t *a, const uint64_t *b, uint64_t *out, int bits);

/* ---- bootstrap ---- */
/* lwe_bootstrap_key_conversion.rs:99-152 : std [n][l][k+1][k+1][N] -> fourier [n][l][k+1][k+1][M] c64 */
void orc_bsk_to_fourier(const orc_params *p, const uint64_t *bsk, double *fbsk, int nthreads);
/* fft64/crypto/ggsw.rs:477-598 : acc += ggsw (x) glwe ; ggsw_f is one Fourier GGSW [l][k+1][k+1][M] */
void orc_add_external_product(const orc_params *p, uint64_t *acc, const double *ggsw_f, const uint64_t *glwe);
/* fft64/crypto/bootstrap.rs:243-344 : acc holds the LUT on entry, rotated GLWE on exit */
void orc_blind_rotate(const orc_params *p, const double *fbsk, const uint64_t *lwe_in, uint64_t *acc);
/* bootstrap.rs:346-380 */
void orc_pbs(const orc_params *p, const double *fbsk, const uint64_t *lwe_in, const uint64_t *lut, uint64_t *lwe_out);
/* shortint/server_key/mod.rs:783-857 (KS then PBS), batched over nthreads like benches/core_crypto/pbs_bench.rs:517-532 */
void orc_ks_pbs_batch(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in,
                      const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads);
/* boolean/engine/bootstrapping.rs:299-344 (PBS then KS) */
void orc_pbs_ks_batch(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in,
                      const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads);
void orc_pbs_batch(const orc_params *p, const double *fbsk, const uint64_t *in, const uint64_t *luts,
                   const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads, int sample_extract);

/* ---- multi-bit PBS (algorithms/lwe_multi_bit_programmable_bootstrapping.rs) ---- */
/* mb_fbsk: [n/G][2^G] Fourier GGSWs (orc_bsk_to_fourier with n := n/G * 2^G) */
void orc_multibit_blind_rotate(const orc_params *p, int grouping, const double *mb_fbsk, const uint64_t *lwe_in, uint64_t *acc);
void orc_multibit_pbs(const orc_params *p, int grouping, const double *mb_fbsk, const uint64_t *lwe_in, const uint64_t *lut, uint64_t *lwe_out);
void orc_multibit_batch(const orc_params *p, int grouping, int mode, const uint64_t *ksk, const double *mb_fbsk, const uint64_t *in,
                        const uint64_t *luts, const uint32_t *lut_idx, uint64_t *out, int batch, int nthreads);

/* ---- gadget extras of the fork (gadget/engine/bootstrapping.rs) ---- */
/* algorithms/lwe_packing_keyswitch.rs:100-187 ; pksk [kN][l_ks][(k+1)N] */
void orc_packing_keyswitch(const orc_params *p, const uint64_t *pksk, const uint64_t *in, uint64_t *out_glwe);
/* mvb_bootstrap post-processing :596-611 : sample_extract_0(glwe * poly) */
void orc_glwe_mul_plain_extract(const orc_params *p, const uint64_t *glwe, const uint64_t *poly, uint64_t *lwe_out);
/* pack_into_new_accumulator :690-773 */
void orc_pack_into_accumulator(const orc_params *p, const uint64_t *pksk, const uint64_t *lwes, int count, int pmod, uint64_t *out_glwe);

/* extract_bits, fft64/crypto/wop_pbs/mod.rs:66-225 ; out [nbits][n+1], most significant extracted bit first */
void orc_extract_bits(const orc_params *p, const uint64_t *ksk, const double *fbsk, const uint64_t *in, int delta_log, int nbits,
                      uint64_t *out);

/* private functional packing keyswitch, algorithms/lwe_private_functional_packing_keyswitch.rs:8-89 ; key [kN+1][level 1..L][(k+1)N] */
void orc_pfpks(const orc_params *p, int base_log, int level, const uint64_t *key, const uint64_t *in, uint64_t *out_glwe);
/* circuit_bootstrap_boolean, wop_pbs/mod.rs:243-344,369-444 ; ggsw_out [cbs_level][k+1][(k+1)N] */
void orc_circuit_bootstrap_boolean(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                                   int cbs_base_log, int cbs_level, const uint64_t *in, int delta_log, uint64_t *ggsw_out);
/* circuit_bootstrap_boolean_vertical_packing, :647-853, one-polynomial look-up tables ; out [n_out][kN+1] */
void orc_cbs_vertical_packing(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                              int cbs_base_log, int cbs_level, const uint64_t *bits_in, int nbits, const uint64_t *luts, int n_out,
                              uint64_t *out);
/* the same with tables of polys_per_lut = 2^t polynomials: CMUX tree (:468-592) over the t most significant bits first */
void orc_cbs_vertical_packing_tree(const orc_params *p, const double *fbsk, const uint64_t *pfpksk_list, int pf_base_log, int pf_level,
                                   int cbs_base_log, int cbs_level, const uint64_t *bits_in, int nbits, const uint64_t *luts, int n_out,
                                   int polys_per_lut, uint64_t *out);
/* generate_circuit_bootstrap_lwe_pfpksk_list, algorithms/lwe_wopbs.rs:80-160 ; out [k+1][kN+1][level 1..L][(k+1)N] */
void orc_gen_cbs_pfpksk_list(const orc_params *p, int base_log, int level, const uint64_t *glwe_key, uint64_t seed, uint64_t *out,
                             int nthreads);

/* ---- LUTs and gate glue ---- */
/* shortint/engine/mod.rs:72-128 ; f_table has p = msg_mod*carry_mod entries */
void orc_fill_accumulator(const orc_params *p, const uint64_t *f_table, int pmod, uint64_t *glwe_out);
/* boolean/engine/bootstrapping.rs:59-60 */
void orc_boolean_lut(const orc_params *p, uint64_t *glwe_out);
/* boolean/engine/mod.rs:607-798 linear combinations; gate: 0 AND 1 NAND 2 OR 3 NOR 4 XOR 5 XNOR */
void orc_boolean_gate_prologue(int gate, int len, const uint64_t *a, const uint64_t *b, uint64_t *out);

/* ---- key / ciphertext synthesis (oracle-only; algorithms/ encryption and key generation files) ---- */
void orc_gen_binary_key(uint64_t seed, int len, uint64_t *key);
void orc_lwe_encrypt(int dim, int bits, const uint64_t *key, uint64_t plaintext, double std, uint64_t seed, uint64_t *ct);
uint64_t orc_lwe_phase(int dim, int bits, const uint64_t *key, const uint64_t *ct);
void orc_gen_ksk(const orc_params *p, const uint64_t *big_key, const uint64_t *small_key, uint64_t seed, uint64_t *ksk, int nthreads);
void orc_gen_bsk(const orc_params *p, const uint64_t *small_key, const uint64_t *glwe_key, uint64_t seed, uint64_t *bsk, int nthreads);
/* lwe_multi_bit_bootstrap_key_generation.rs:87-172: [n/G][2^G][l][k+1][k+1][N] */
void orc_gen_multibit_bsk(const orc_params *p, int grouping, const uint64_t *small_key, const uint64_t *glwe_key, uint64_t seed,
                          uint64_t *bsk, int nthreads);
/* lwe_packing_keyswitch_key_generation.rs:60-150 */
void orc_gen_pksk(const orc_params *p, const uint64_t *big_key, const uint64_t *glwe_key, uint64_t seed, uint64_t *pksk, int nthreads);
/* exact negacyclic product of a torus poly by a binary poly (limb-split FFT), used by GLWE encryption */
void orc_negacyclic_mul_binary(int N, const uint64_t *a, const uint64_t *s_binary, uint64_t *out, int bits);

#ifdef __cplusplus
}
#endif
#endif

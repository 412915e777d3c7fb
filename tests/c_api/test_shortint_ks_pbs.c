/* C client of the drop-in boundary (include/b200tfhe.h), in the style of the reference's tfhe/c_api_tests/test_shortint_*.c:
 * create a context, load a keyswitch key and a bootstrap key, run shortint's keyswitch + programmable bootstrap on a list of
 * ciphertexts through the host-pointer entry points, check every decrypted result, exercise the error convention, free
 * everything.  Keys, encryption and decryption come from the CPU oracle (tests only; the library under test never sees it).
 *
 * PARAM_MESSAGE_2_CARRY_2_KS_PBS dimensions with a short LWE dimension so that the key generation takes a second
 * (the reference's TEST_PARAMS_4_BITS_NATIVE_U64 shape, core_crypto/algorithms/test/mod.rs:55-72).
 *
 * build + run: tests/test_c_api.py */
#include <assert.h>
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200tfhe.h"
#include "tfhe_oracle.h"

#define CHECK(call)                                                                                  \
    do {                                                                                             \
        int rc_ = (call);                                                                            \
        if (rc_ != 0) {                                                                              \
            fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, ctx ? b200tfhe_last_error(ctx) : ""); \
            return 1;                                                                                \
        }                                                                                            \
    } while (0)

int main(int argc, char **argv) {
    const int compute = argc < 2 || strcmp(argv[1], "--link-only") != 0;
    const int n = 64, k = 1, N = 2048, pmod = 16, batch = 40;
    orc_params op = {n, k, N, 23, 1, 3, 5, 64, 1e-9, 0.00000000000000029403601535432533};
    b200tfhe_params p = {n, k, N, 23, 1, 3, 5, 64};
    const int big = k * N, glwe_len = (k + 1) * N;
    b200tfhe_context *ctx = NULL;

    if (!compute) {   /* no GPU on this machine: the program links against every symbol it names, nothing runs */
        printf("link ok: %p %p %p\n", (void *)&b200tfhe_context_create, (void *)&b200tfhe_ks_pbs_batch_u64, (void *)&b200tfhe_keyswitch_batch_u64);
        return 0;
    }

    /* keys (oracle) */
    uint64_t *small_key = calloc(n, 8), *glwe_key = calloc(big, 8);
    uint64_t *ksk = calloc((size_t)big * 5 * (n + 1), 8), *bsk = calloc((size_t)n * 1 * 4 * N, 8);
    orc_gen_binary_key(0xB2000001ull, n, small_key);
    orc_gen_binary_key(0xB2000002ull, big, glwe_key);
    orc_gen_ksk(&op, glwe_key, small_key, 0xB2000003ull, ksk, 4);
    orc_gen_bsk(&op, small_key, glwe_key, 0xB2000004ull, bsk, 4);

    /* accumulator of f(x) = x*x + 1 mod 16 (shortint generate_lookup_table) */
    uint64_t table[16], *lut = calloc(glwe_len, 8);
    for (int x = 0; x < pmod; x++) table[x] = (uint64_t)((x * x + 1) % pmod);
    orc_fill_accumulator(&op, table, pmod, lut);

    /* inputs: every message, encrypted under the big key; delta = 2^63 / 16 */
    uint64_t *in = calloc((size_t)batch * (big + 1), 8), *out = calloc((size_t)batch * (big + 1), 8), *mid = calloc((size_t)batch * (n + 1), 8);
    for (int i = 0; i < batch; i++)
        orc_lwe_encrypt(big, 64, glwe_key, (uint64_t)(i % pmod) << 59, op.glwe_std, 0xC0DE + i, in + (size_t)i * (big + 1));

    CHECK(b200tfhe_context_create(0, &ctx));
    b200tfhe_ksk *dksk = NULL;
    b200tfhe_bsk *dbsk = NULL;
    CHECK(b200tfhe_ksk_upload_u64(ctx, &p, ksk, &dksk));
    CHECK(b200tfhe_bsk_upload_u64(ctx, &p, bsk, &dbsk));

    /* keyswitch alone is bit-exact against the reference algorithm */
    CHECK(b200tfhe_keyswitch_batch_u64(ctx, dksk, in, mid, batch));
    uint64_t *ref_small = calloc(n + 1, 8);
    for (int i = 0; i < batch; i++) {
        orc_keyswitch(&op, ksk, in + (size_t)i * (big + 1), ref_small);
        if (memcmp(ref_small, mid + (size_t)i * (n + 1), (size_t)(n + 1) * 8) != 0) {
            fprintf(stderr, "keyswitch of ciphertext %d differs from the reference algorithm\n", i);
            return 1;
        }
    }

    /* ServerKey::apply_lookup_table core: keyswitch then programmable bootstrap */
    CHECK(b200tfhe_ks_pbs_batch_u64(ctx, dksk, dbsk, in, out, lut, 1, NULL, batch));
    for (int i = 0; i < batch; i++) {
        const uint64_t phase = orc_lwe_phase(big, 64, glwe_key, out + (size_t)i * (big + 1));
        const int got = (int)(((phase + (1ull << 58)) >> 59) & 15), want = ((i % pmod) * (i % pmod) + 1) % pmod;
        if (got != want) {
            fprintf(stderr, "ciphertext %d: decrypted %d, expected %d\n", i, got, want);
            return 1;
        }
    }

    /* error convention (c_api/utils.rs:3-27): non-zero return, message available, nothing written through a NULLed out-pointer */
    b200tfhe_params bad = p;
    bad.polynomial_size = 1000;
    b200tfhe_bsk *nope = (b200tfhe_bsk *)0x1;
    assert(b200tfhe_bsk_upload_u64(ctx, &bad, bsk, &nope) != 0);
    assert(nope == NULL);
    assert(strlen(b200tfhe_last_error(ctx)) > 0);
    assert(b200tfhe_ks_pbs_batch_u64(ctx, dksk, dbsk, NULL, out, lut, 1, NULL, batch) != 0);

    CHECK(b200tfhe_bsk_destroy(dbsk));
    CHECK(b200tfhe_ksk_destroy(dksk));
    CHECK(b200tfhe_context_destroy(ctx));
    free(small_key); free(glwe_key); free(ksk); free(bsk); free(lut); free(in); free(out); free(mid); free(ref_small);
    printf("ok: %d ciphertexts, keyswitch bit-exact, every bootstrap decrypts to f(m)\n", batch);
    return 0;
}

// multi_gpu.cu -- several GPUs of one node behind ONE handle (SURVEY 8(b) `ctx_create(n_gpus)`, 8(e)).
//
// The path shards by independent ciphertexts (the reference: rayon par_iter over ciphertexts, tfhe/benches/core_crypto/
// pbs_bench.rs:517-532), so a multi-GPU call is: a contiguous static split of the batch index range (remainder to the low
// devices), one host thread + one context (stream, workspace) per GPU, every GPU reading its slice of the caller's input buffer
// and writing its slice of the caller's output buffer at the same indices.  No per-batch collective.  The only inter-GPU traffic is
// the one-time fan-out of the keys: uploaded and converted to the Fourier domain on device 0, then copied device-to-device
// (cudaMemcpyPeerAsync: NVLink / NVSwitch where the topology has it) -- the Fourier layout is identical on every B200.
// Built entirely on the single-GPU entry points of b200tfhe.h.
#include "../../include/b200tfhe.h"
#include <cuda_runtime.h>

#include <string>
#include <thread>
#include <vector>

struct b200tfhe_multi_context {
    std::vector<b200tfhe_context *> ctx;   // ctx[g] on device g
    std::string err;
};
struct b200tfhe_multi_keys {
    b200tfhe_multi_context *m;
    b200tfhe_params p;
    std::vector<b200tfhe_bsk *> bsk;       // may be empty (keyswitch only)
    std::vector<b200tfhe_ksk *> ksk;       // may be empty (bootstrap only)
};

static thread_local std::string g_multi_err;

static int mfail(b200tfhe_multi_context *m, const std::string &msg) {
    if (m) m->err = msg; else g_multi_err = msg;
    return 1;
}

// [begin, end) of the ciphertexts of device g: contiguous, sizes differ by at most one (the same rule as parallel.shard_range)
static void shard(uint32_t batch, int n, int g, uint32_t *begin, uint32_t *end) {
    const uint32_t base = batch / (uint32_t)n, rem = batch % (uint32_t)n;
    *begin = (uint32_t)g * base + ((uint32_t)g < rem ? (uint32_t)g : rem);
    *end = *begin + base + ((uint32_t)g < rem ? 1u : 0u);
}

extern "C" {

const char *b200tfhe_multi_last_error(const b200tfhe_multi_context *m) { return m ? m->err.c_str() : g_multi_err.c_str(); }

int b200tfhe_multi_context_create(int n_gpus, b200tfhe_multi_context **out) {
    if (!out) return mfail(nullptr, "multi_context_create: null out pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return mfail(nullptr, std::string("multi_context_create: no CUDA device (") + cudaGetErrorString(e) + ")");
    if (n_gpus <= 0) n_gpus = count;
    if (n_gpus > count) return mfail(nullptr, "multi_context_create: asked for " + std::to_string(n_gpus) + " GPUs, " + std::to_string(count) + " visible");
    b200tfhe_multi_context *m = new b200tfhe_multi_context();
    for (int g = 0; g < n_gpus; g++) {
        b200tfhe_context *c = nullptr;
        if (b200tfhe_context_create(g, &c)) {
            std::string why = b200tfhe_last_error(nullptr);
            for (auto *x : m->ctx) b200tfhe_context_destroy(x);
            delete m;
            return mfail(nullptr, "multi_context_create: device " + std::to_string(g) + ": " + why);
        }
        m->ctx.push_back(c);
    }
    // direct peer copies for the key fan-out (harmless if already enabled / unsupported: the copy then stages through the host)
    for (int g = 1; g < n_gpus; g++) {
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, g, 0) == cudaSuccess && can) {
            cudaSetDevice(g);
            cudaDeviceEnablePeerAccess(0, 0);
        }
    }
    cudaGetLastError();
    cudaSetDevice(0);
    *out = m;
    return 0;
}

int b200tfhe_multi_context_destroy(b200tfhe_multi_context *m) {
    if (!m) return 0;
    for (auto *c : m->ctx) b200tfhe_context_destroy(c);
    delete m;
    return 0;
}

int b200tfhe_multi_context_gpus(const b200tfhe_multi_context *m) { return m ? (int)m->ctx.size() : 0; }

b200tfhe_context *b200tfhe_multi_context_get(b200tfhe_multi_context *m, int gpu) {
    return (m && gpu >= 0 && gpu < (int)m->ctx.size()) ? m->ctx[gpu] : nullptr;
}

int b200tfhe_multi_keys_destroy(b200tfhe_multi_keys *k) {
    if (!k) return 0;
    for (auto *b : k->bsk) b200tfhe_bsk_destroy(b);
    for (auto *s : k->ksk) b200tfhe_ksk_destroy(s);
    delete k;
    return 0;
}

static int multi_keys_upload(b200tfhe_multi_context *m, const b200tfhe_params *p, const void *std_bsk, const void *ksk, int bits,
                             b200tfhe_multi_keys **out) {
    if (!m || !p || !out) return mfail(m, "multi_keys_upload: null argument");
    *out = nullptr;
    if (!std_bsk && !ksk) return mfail(m, "multi_keys_upload: neither a bootstrap key nor a keyswitch key given");
    if (p->torus_bits != bits) return mfail(m, "multi_keys_upload: torus_bits does not match the entry point");
    const int n = (int)m->ctx.size();
    b200tfhe_multi_keys *k = new b200tfhe_multi_keys{m, *p, {}, {}};
    auto bail = [&](const std::string &why) {
        b200tfhe_multi_keys_destroy(k);
        return mfail(m, why);
    };
    // device 0: upload (+ conversion to the Fourier domain / tensor-core key planes)
    if (std_bsk) {
        b200tfhe_bsk *b0 = nullptr;
        int rc = bits == 64 ? b200tfhe_bsk_upload_u64(m->ctx[0], p, (const uint64_t *)std_bsk, &b0) : b200tfhe_bsk_upload_u32(m->ctx[0], p, (const uint32_t *)std_bsk, &b0);
        if (rc) return bail(std::string("multi_keys_upload: bootstrap key: ") + b200tfhe_last_error(m->ctx[0]));
        k->bsk.push_back(b0);
    }
    if (ksk) {
        b200tfhe_ksk *s0 = nullptr;
        int rc = bits == 64 ? b200tfhe_ksk_upload_u64(m->ctx[0], p, (const uint64_t *)ksk, &s0) : b200tfhe_ksk_upload_u32(m->ctx[0], p, (const uint32_t *)ksk, &s0);
        if (rc) return bail(std::string("multi_keys_upload: keyswitch key: ") + b200tfhe_last_error(m->ctx[0]));
        k->ksk.push_back(s0);
    }
    // fan-out: device-to-device copies of the device-layout keys, then the per-device finalisation of the keyswitch key
    for (int g = 1; g < n; g++) {
        if (std_bsk) {
            b200tfhe_bsk *bg = nullptr;
            if (b200tfhe_bsk_alloc(m->ctx[g], p, &bg)) return bail(std::string("multi_keys_upload: bsk_alloc: ") + b200tfhe_last_error(m->ctx[g]));
            k->bsk.push_back(bg);
            void *src, *dst; size_t bytes, bytes_g;
            b200tfhe_bsk_device_buffer(k->bsk[0], &src, &bytes);
            b200tfhe_bsk_device_buffer(bg, &dst, &bytes_g);
            cudaError_t e = cudaMemcpyPeerAsync(dst, g, src, 0, bytes, (cudaStream_t)b200tfhe_context_stream(m->ctx[g]));
            if (e != cudaSuccess) return bail(std::string("multi_keys_upload: peer copy of the bootstrap key: ") + cudaGetErrorString(e));
        }
        if (ksk) {
            b200tfhe_ksk *sg = nullptr;
            if (b200tfhe_ksk_alloc(m->ctx[g], p, &sg)) return bail(std::string("multi_keys_upload: ksk_alloc: ") + b200tfhe_last_error(m->ctx[g]));
            k->ksk.push_back(sg);
            void *src, *dst; size_t bytes, bytes_g;
            b200tfhe_ksk_device_buffer(k->ksk[0], &src, &bytes);
            b200tfhe_ksk_device_buffer(sg, &dst, &bytes_g);
            cudaError_t e = cudaMemcpyPeerAsync(dst, g, src, 0, bytes, (cudaStream_t)b200tfhe_context_stream(m->ctx[g]));
            if (e != cudaSuccess) return bail(std::string("multi_keys_upload: peer copy of the keyswitch key: ") + cudaGetErrorString(e));
        }
    }
    for (int g = 1; g < n; g++) {
        if (b200tfhe_synchronize(m->ctx[g])) return bail(std::string("multi_keys_upload: ") + b200tfhe_last_error(m->ctx[g]));
        if (ksk && b200tfhe_ksk_finalize(m->ctx[g], k->ksk[g])) return bail(std::string("multi_keys_upload: ksk_finalize: ") + b200tfhe_last_error(m->ctx[g]));
    }
    cudaSetDevice(0);
    *out = k;
    return 0;
}

int b200tfhe_multi_keys_upload_u64(b200tfhe_multi_context *m, const b200tfhe_params *p, const uint64_t *std_bsk, const uint64_t *ksk,
                                   b200tfhe_multi_keys **out) {
    return multi_keys_upload(m, p, std_bsk, ksk, 64, out);
}
int b200tfhe_multi_keys_upload_u32(b200tfhe_multi_context *m, const b200tfhe_params *p, const uint32_t *std_bsk, const uint32_t *ksk,
                                   b200tfhe_multi_keys **out) {
    return multi_keys_upload(m, p, std_bsk, ksk, 32, out);
}

}  // extern "C"

// op: 0 keyswitch, 1 pbs, 2 ks_pbs, 3 pbs_ks
static int multi_op(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, int op, int bits, const void *in, void *out, const void *luts,
                    uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch) {
    if (!m || !k || k->m != m) return mfail(m, "multi op: bad handle");
    const bool need_ksk = op != 1, need_bsk = op != 0;
    if ((need_ksk && k->ksk.empty()) || (need_bsk && k->bsk.empty())) return mfail(m, "multi op: the key set lacks the key this operation needs");
    if (k->p.torus_bits != bits) return mfail(m, "multi op: torus width of the keys does not match the entry point");
    if (batch == 0) return 0;
    if (!in || !out) return mfail(m, "multi op: null ciphertext buffer");
    const int n = (int)m->ctx.size();
    const size_t s = (size_t)bits / 8;
    const size_t big = (size_t)k->p.glwe_dimension * k->p.polynomial_size + 1, small = (size_t)k->p.lwe_dimension + 1;
    const size_t in_len = (op == 0 || op == 2) ? big : small, out_len = (op == 0 || op == 3) ? small : big;
    std::vector<int> rc(n, 0);
    std::vector<std::thread> th;
    for (int g = 0; g < n; g++) {
        uint32_t b0, b1;
        shard(batch, n, g, &b0, &b1);
        if (b1 == b0) continue;
        th.emplace_back([=, &rc]() {
            const char *pi = (const char *)in + (size_t)b0 * in_len * s;
            char *po = (char *)out + (size_t)b0 * out_len * s;
            const uint32_t *li = lut_idx ? lut_idx + b0 : nullptr;
            const uint32_t cb = b1 - b0;
            b200tfhe_context *c = m->ctx[g];
            const b200tfhe_ksk *ks = need_ksk ? k->ksk[g] : nullptr;
            const b200tfhe_bsk *bs = need_bsk ? k->bsk[g] : nullptr;
            if (bits == 64) {
                const uint64_t *i64 = (const uint64_t *)pi; uint64_t *o64 = (uint64_t *)po; const uint64_t *l64 = (const uint64_t *)luts;
                rc[g] = op == 0 ? b200tfhe_keyswitch_batch_u64(c, ks, i64, o64, cb)
                      : op == 1 ? b200tfhe_pbs_batch_u64(c, bs, i64, o64, l64, n_luts, li, cb)
                      : op == 2 ? b200tfhe_ks_pbs_batch_u64(c, ks, bs, i64, o64, l64, n_luts, li, cb)
                                : b200tfhe_pbs_ks_batch_u64(c, ks, bs, i64, o64, l64, n_luts, li, cb);
            } else {
                const uint32_t *i32 = (const uint32_t *)pi; uint32_t *o32 = (uint32_t *)po; const uint32_t *l32 = (const uint32_t *)luts;
                rc[g] = op == 0 ? b200tfhe_keyswitch_batch_u32(c, ks, i32, o32, cb)
                      : op == 1 ? b200tfhe_pbs_batch_u32(c, bs, i32, o32, l32, n_luts, li, cb)
                      : op == 2 ? b200tfhe_ks_pbs_batch_u32(c, ks, bs, i32, o32, l32, n_luts, li, cb)
                                : b200tfhe_pbs_ks_batch_u32(c, ks, bs, i32, o32, l32, n_luts, li, cb);
            }
        });
    }
    for (auto &t : th) t.join();
    for (int g = 0; g < n; g++)
        if (rc[g]) return mfail(m, "multi op: device " + std::to_string(g) + ": " + b200tfhe_last_error(m->ctx[g]));
    return 0;
}

extern "C" {

#define MULTI_OPS(SUF, T, BITS)                                                                                                              \
    int b200tfhe_multi_keyswitch_batch_##SUF(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const T *in, T *out, uint32_t b) {     \
        return multi_op(m, k, 0, BITS, in, out, nullptr, 0, nullptr, b);                                                                     \
    }                                                                                                                                        \
    int b200tfhe_multi_pbs_batch_##SUF(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const T *in, T *out, const T *l,             \
                                       uint32_t nl, const uint32_t *li, uint32_t b) {                                                        \
        return multi_op(m, k, 1, BITS, in, out, l, nl, li, b);                                                                               \
    }                                                                                                                                        \
    int b200tfhe_multi_ks_pbs_batch_##SUF(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const T *in, T *out, const T *l,          \
                                          uint32_t nl, const uint32_t *li, uint32_t b) {                                                     \
        return multi_op(m, k, 2, BITS, in, out, l, nl, li, b);                                                                               \
    }                                                                                                                                        \
    int b200tfhe_multi_pbs_ks_batch_##SUF(b200tfhe_multi_context *m, const b200tfhe_multi_keys *k, const T *in, T *out, const T *l,          \
                                          uint32_t nl, const uint32_t *li, uint32_t b) {                                                     \
        return multi_op(m, k, 3, BITS, in, out, l, nl, li, b);                                                                               \
    }
MULTI_OPS(u64, uint64_t, 64)
MULTI_OPS(u32, uint32_t, 32)

}  // extern "C"

// api.cu -- the extern "C" boundary declared in include/b200tfhe.h.
// Host-side logic only: argument validation (the reference's assert!s, lwe_keyswitch.rs:106-141,
// lwe_programmable_bootstrapping.rs:1083-1097), workspace management, H2D/D2H staging and kernel launches.
#include "../../include/b200tfhe.h"
#include "kernels.h"
#include "twiddles.hpp"
#include <mutex>
#include <string>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>

using namespace b200tfhe;

static thread_local std::string g_global_err = "";

struct Workspace {
    enum { IN = 0, OUT, MID, LUTS, IDX, A, B, DIG, C, D, MBK, NSLOTS };
    void *ptr[NSLOTS] = {};
    size_t cap[NSLOTS] = {};
    cudaError_t get(int slot, size_t bytes, void **out) {
        if (bytes == 0) bytes = 16;
        if (cap[slot] < bytes) {
            if (ptr[slot]) cudaFree(ptr[slot]);
            ptr[slot] = nullptr;
            cap[slot] = 0;
            cudaError_t e = cudaMalloc(&ptr[slot], bytes);
            if (e != cudaSuccess) return e;
            cap[slot] = bytes;
        }
        *out = ptr[slot];
        return cudaSuccess;
    }
    void release() {
        for (int i = 0; i < NSLOTS; i++) {
            if (ptr[i]) cudaFree(ptr[i]);
            ptr[i] = nullptr;
            cap[i] = 0;
        }
    }
};

struct b200tfhe_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    std::string err;
    void *tw16 = nullptr, *tw32 = nullptr;  // TwTables<R> on the device
    void *twbig = nullptr;                  // BigTables (N = 8192)
    Workspace ws;
    uint64_t launches = 0;
    // host-pointer calls on large batches are cut into chunks: H2D of chunk i+1 and D2H of chunk i-1 overlap the kernels of chunk i
    enum { MAX_CHUNKS = 8 };
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    cudaEvent_t ev_in[MAX_CHUNKS] = {}, ev_done[MAX_CHUNKS] = {}, ev_start = nullptr;
    // tuning / test knobs (b200tfhe_context_set_option); the defaults are what production uses
    int opt_br_kernel = BR_AUTO;     // which blind-rotation kernel (kernels.h: BR_AUTO .. BR_WS)
    int opt_lat_max = 2 * 148;       // largest batch on the one-ciphertext-per-SM latency kernel
    int opt_lat_cluster_max = 74;    // largest batch that gives every ciphertext a two-CTA cluster (two SMs) on the latency kernel
    int opt_mb_lat_max = 148;        // multi-bit keys: largest batch on the two-kernel latency path (16-24 MB of scratch per ciphertext)
    int opt_ks_tensor_core = 1;      // 0: CUDA-core keyswitch even where the tensor-core planes exist
    int opt_ks_tc_min = 9;           // smallest batch on the tensor-core keyswitch (up to 8: the split-row kernel, which leaves L2 to the bootstrap key)
    uint32_t opt_chunk_min = 6 * 592;   // minimum chunk of the H2D / compute / D2H pipeline of host-pointer calls (0: off)
};
struct b200tfhe_bsk {
    b200tfhe_context *ctx;
    Params p;
    void *d_fourier;
    size_t bytes;
    int grouping = 0;    // 0: classic key (one GGSW per mask element); G >= 2: multi-bit key, 2^G GGSWs per G elements
    int device = 0;      // copied from the context: *_destroy must not dereference a context that may already be gone
};
struct b200tfhe_ksk {
    b200tfhe_context *ctx;
    Params p;
    void *d_ksk, *d_colsum;
    size_t bytes;
    bool finalized;
    KSTensorKey tk;      // byte-plane copy of the key for the tensor-core path (d_planes == nullptr: not available)
    int body_col = -1;   // output column receiving the input body (-1: the last one); k*N for a packing keyswitch key
    int sub = 1, sub_bits = 0;   // base_log > 7: every digit split exactly into `sub` int8 sub-digits (KSLaunch::sub), tensor-core path only
    int device = 0;
};
// LwePackingKeyswitchKey: the same integer GEMM with (k+1)N output columns and the body on coefficient 0 of the body polynomial
struct b200tfhe_pksk {
    b200tfhe_ksk *inner;
    Params p;
};

static int fail(b200tfhe_context *ctx, const std::string &msg) {
    if (ctx) ctx->err = msg; else g_global_err = msg;
    return 1;
}
static int fail_cuda(b200tfhe_context *ctx, const char *what, cudaError_t e) {
    return fail(ctx, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                                   \
    do {                                                           \
        cudaError_t e_ = (call);                                   \
        if (e_ != cudaSuccess) return fail_cuda(ctx, #call, e_);   \
    } while (0)
#define REQUIRE(cond, msg)                       \
    do {                                         \
        if (!(cond)) return fail(ctx, (msg));    \
    } while (0)

// Host-pointer entry points enqueue copies from / into the CALLER's buffers: on any early return after the first enqueue the streams
// are drained first, so the GPU never touches a buffer the caller may already have freed after seeing the error.
struct DrainOnExit {
    b200tfhe_context *ctx;
    bool armed = true;
    explicit DrainOnExit(b200tfhe_context *c) : ctx(c) {}
    ~DrainOnExit() {
        if (!armed) return;
        cudaStreamSynchronize(ctx->stream);
        if (ctx->s_h2d) cudaStreamSynchronize(ctx->s_h2d);
        if (ctx->s_d2h) cudaStreamSynchronize(ctx->s_d2h);
    }
};

static Params to_params(const b200tfhe_params *p) {
    Params q;
    q.n = p->lwe_dimension; q.k = p->glwe_dimension; q.N = p->polynomial_size;
    q.pbs_base_log = p->pbs_base_log; q.pbs_level = p->pbs_level;
    q.ks_base_log = p->ks_base_log; q.ks_level = p->ks_level; q.bits = p->torus_bits;
    return q;
}
static bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }
static size_t sz(const Params &p) { return (size_t)p.bits / 8; }
static size_t big_len(const Params &p) { return (size_t)p.k * p.N + 1; }
static size_t small_len(const Params &p) { return (size_t)p.n + 1; }
static size_t glwe_len(const Params &p) { return (size_t)(p.k + 1) * p.N; }

static const void *tw_fwd(const b200tfhe_context *ctx, int R) {
    if (R == 256) return ctx->twbig;
    return R == 32 ? (const void *)((TwTables<32> *)ctx->tw32)->fwd : (const void *)((TwTables<16> *)ctx->tw16)->fwd;
}
static const void *tw_inv(const b200tfhe_context *ctx, int R) {
    if (R == 256) return ctx->twbig;
    return R == 32 ? (const void *)((TwTables<32> *)ctx->tw32)->inv : (const void *)((TwTables<16> *)ctx->tw16)->inv;
}

extern "C" {

const char *b200tfhe_last_error(const b200tfhe_context *ctx) { return ctx ? ctx->err.c_str() : g_global_err.c_str(); }

int b200tfhe_context_create(int device, b200tfhe_context **out) {
    b200tfhe_context *ctx = nullptr;
    if (!out) return fail(nullptr, "context_create: null out pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, std::string("context_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU fallback");
    if (device < 0 || device >= count) return fail(nullptr, "context_create: device index out of range");
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail_cuda(nullptr, "cudaSetDevice", e);
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return fail_cuda(nullptr, "cudaGetDeviceProperties", e);
    if (prop.major != 10) return fail(nullptr, std::string("context_create: built for sm_100a only, device is ") + prop.name);
    ctx = new b200tfhe_context();
    ctx->device = device;
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete ctx; return fail_cuda(nullptr, "cudaStreamCreate", e); }
    {
        static TwTables<16> t16;
        static TwTables<32> t32;
        static BigTables tbig;
        static std::once_flag once;
        std::call_once(once, [] { fill_tw_tables(t16); fill_tw_tables(t32); fill_big_tables(tbig); });
        if ((e = cudaMalloc(&ctx->tw16, sizeof(t16))) != cudaSuccess || (e = cudaMalloc(&ctx->tw32, sizeof(t32))) != cudaSuccess ||
            (e = cudaMalloc(&ctx->twbig, sizeof(tbig))) != cudaSuccess ||
            (e = cudaMemcpy(ctx->twbig, &tbig, sizeof(tbig), cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(ctx->tw16, &t16, sizeof(t16), cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(ctx->tw32, &t32, sizeof(t32), cudaMemcpyHostToDevice)) != cudaSuccess) {
            delete ctx;
            return fail_cuda(nullptr, "twiddle upload", e);
        }
    }
    *out = ctx;
    return 0;
}

int b200tfhe_context_destroy(b200tfhe_context *ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->ws.release();
    cudaFree(ctx->tw16);
    cudaFree(ctx->tw32);
    cudaFree(ctx->twbig);
    cudaStreamDestroy(ctx->stream);
    if (ctx->s_h2d) cudaStreamDestroy(ctx->s_h2d);
    if (ctx->s_d2h) cudaStreamDestroy(ctx->s_d2h);
    for (int i = 0; i < b200tfhe_context::MAX_CHUNKS; i++) {
        if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
    }
    if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
    delete ctx;
    return 0;
}

int b200tfhe_synchronize(b200tfhe_context *ctx) {
    REQUIRE(ctx, "synchronize: null context");
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
void *b200tfhe_context_stream(b200tfhe_context *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
uint64_t b200tfhe_launch_count(const b200tfhe_context *ctx) { return ctx ? ctx->launches : 0; }

int b200tfhe_context_set_option(b200tfhe_context *ctx, const char *key, int64_t value) {
    REQUIRE(ctx && key, "context_set_option: null argument");
    std::lock_guard<std::mutex> g(ctx->mu);
    const std::string k(key);
    if (k == "br_kernel") {
        REQUIRE(value >= BR_AUTO && value <= BR_PAIR, "context_set_option: br_kernel must be 0 (auto) .. 7");
        ctx->opt_br_kernel = (int)value;
    } else if (k == "x16_variant") {
        br_x16_set_variant((int)value);
    } else if (k == "lat_max") {
        REQUIRE(value >= 0 && value <= (1 << 20), "context_set_option: lat_max out of range");
        ctx->opt_lat_max = (int)value;
    } else if (k == "lat_cluster_max") {
        REQUIRE(value >= 0 && value <= 4096, "context_set_option: lat_cluster_max out of range (0 .. 4096)");
        ctx->opt_lat_cluster_max = (int)value;
    } else if (k == "mb_lat_max") {
        REQUIRE(value >= 0 && value <= 4096, "context_set_option: mb_lat_max out of range (0 .. 4096)");
        ctx->opt_mb_lat_max = (int)value;
    } else if (k == "ks_tensor_core") {
        ctx->opt_ks_tensor_core = value != 0;
    } else if (k == "ks_tc_min") {
        REQUIRE(value >= 0, "context_set_option: ks_tc_min must be >= 0");
        ctx->opt_ks_tc_min = (int)value;
    } else if (k == "chunk_min") {
        REQUIRE(value >= 0 && value <= 0xFFFFFFFFll, "context_set_option: chunk_min out of range");
        ctx->opt_chunk_min = (uint32_t)value;
    } else {
        return fail(ctx, "context_set_option: unknown option \"" + k + "\"");
    }
    return 0;
}

int b200tfhe_params_supported(const b200tfhe_params *p) {
    if (!p) return 0;
    return br_supported(to_params(p)) != 0;
}

// ---- keys ------------------------------------------------------------------------------------------------
static int check_pbs_params(b200tfhe_context *ctx, const Params &p) {
    REQUIRE(p.bits == 64 || p.bits == 32, "only native 2^64 / 2^32 ciphertext moduli are supported");
    REQUIRE(p.n > 0 && p.k > 0 && is_pow2(p.N), "invalid dimensions");
    REQUIRE(p.pbs_level > 0 && p.pbs_base_log > 0 && p.pbs_level * p.pbs_base_log < p.bits, "invalid PBS decomposition");
    REQUIRE(br_supported(p) != 0, "blind rotation is not built for this (N, k, level, base_log, bits) set");
    return 0;
}
static int check_ks_params(b200tfhe_context *ctx, const Params &p) {
    REQUIRE(p.bits == 64 || p.bits == 32, "only native 2^64 / 2^32 ciphertext moduli are supported");
    REQUIRE(p.n > 0 && p.k > 0 && p.N > 0, "invalid dimensions");
    // base_log <= 7: int8 digits as they are; up to 28: digits split exactly into int8 sub-digits for the tensor-core GEMM
    REQUIRE(p.ks_level > 0 && p.ks_base_log > 0 && p.ks_base_log <= 28 && p.ks_level * p.ks_base_log < p.bits,
            "invalid keyswitch decomposition (base_log must be <= 28)");
    return 0;
}
// number of GGSW ciphertexts of a bootstrap key: n (classic) or n/G * 2^G (entities/lwe_multi_bit_bootstrap_key.rs)
static size_t bsk_ggsw_count(const Params &p, int grouping) {
    return grouping ? (size_t)(p.n / grouping) << grouping : (size_t)p.n;
}
static size_t bsk_polys(const Params &p, int grouping) { return bsk_ggsw_count(p, grouping) * p.pbs_level * (p.k + 1) * (p.k + 1); }
static int bsk_radix(const Params &p, int grouping) { return grouping ? br_mb_supported(p, grouping) : br_supported(p); }
static int check_bsk_params(b200tfhe_context *ctx, const Params &p, int grouping) {
    if (!grouping) return check_pbs_params(ctx, p);
    REQUIRE(p.bits == 64 || p.bits == 32, "only native 2^64 / 2^32 ciphertext moduli are supported");
    REQUIRE(p.n > 0 && p.k > 0 && is_pow2(p.N), "invalid dimensions");
    REQUIRE(p.pbs_level > 0 && p.pbs_base_log > 0 && p.pbs_level * p.pbs_base_log < p.bits, "invalid PBS decomposition");
    // the reference asserts the same divisibility when it builds the key (lwe_multi_bit_bootstrap_key.rs, from_container)
    REQUIRE(grouping >= 2 && grouping <= 4 && p.n % grouping == 0, "grouping factor must be 2..4 and divide the LWE dimension");
    REQUIRE(br_mb_supported(p, grouping) != 0, "multi-bit blind rotation is not built for this (N, k, level, base_log, bits, grouping) set");
    return 0;
}

static int bsk_alloc(b200tfhe_context *ctx, const b200tfhe_params *pp, int grouping, b200tfhe_bsk **out) {
    REQUIRE(ctx && pp && out, "bsk_alloc: null argument");
    *out = nullptr;
    Params p = to_params(pp);
    if (check_bsk_params(ctx, p, grouping)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    b200tfhe_bsk *k = new b200tfhe_bsk{ctx, p, nullptr, bsk_polys(p, grouping) * (p.N / 2) * sizeof(cplx), grouping};
    k->device = ctx->device;
    cudaError_t e = cudaMalloc(&k->d_fourier, k->bytes);
    if (e != cudaSuccess) { delete k; return fail_cuda(ctx, "cudaMalloc(bsk)", e); }
    *out = k;
    return 0;
}

static int bsk_convert_dev(b200tfhe_context *ctx, const b200tfhe_params *pp, int grouping, const void *d_std, b200tfhe_bsk **out) {
    REQUIRE(ctx && pp && out && d_std, "bsk_convert_dev: null argument");
    if (bsk_alloc(ctx, pp, grouping, out)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    Params p = (*out)->p;
    int R = bsk_radix(p, grouping);
    int rc = grouping ? bsk_convert_polys_launch(R, p.bits, bsk_polys(p, grouping), d_std, (*out)->d_fourier, tw_fwd(ctx, R), ctx->stream)
                      : bsk_convert_launch(p, d_std, (*out)->d_fourier, tw_fwd(ctx, R), ctx->stream);
    ctx->launches++;
    if (rc) { b200tfhe_bsk_destroy(*out); *out = nullptr; return fail(ctx, "bsk_convert launch failed: " + std::to_string(rc)); }
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { b200tfhe_bsk_destroy(*out); *out = nullptr; return fail_cuda(ctx, "bsk_convert", e); }
    return 0;
}

static int bsk_upload(b200tfhe_context *ctx, const b200tfhe_params *pp, int grouping, const void *std_bsk, int bits, b200tfhe_bsk **out) {
    REQUIRE(ctx && pp && out && std_bsk, "bsk_upload: null argument");
    *out = nullptr;
    REQUIRE(pp->torus_bits == bits, "bsk_upload: torus_bits does not match the entry point");
    Params p = to_params(pp);
    if (check_bsk_params(ctx, p, grouping)) return 1;
    void *d_std = nullptr;
    size_t bytes = bsk_polys(p, grouping) * p.N * sz(p);
    CK(cudaSetDevice(ctx->device));
    CK(cudaMalloc(&d_std, bytes));
    cudaError_t e = cudaMemcpy(d_std, std_bsk, bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { cudaFree(d_std); return fail_cuda(ctx, "cudaMemcpy(bsk)", e); }
    int rc = bsk_convert_dev(ctx, pp, grouping, d_std, out);
    cudaFree(d_std);
    return rc;
}
int b200tfhe_bsk_alloc(b200tfhe_context *ctx, const b200tfhe_params *p, b200tfhe_bsk **out) { return bsk_alloc(ctx, p, 0, out); }
int b200tfhe_bsk_convert_dev(b200tfhe_context *ctx, const b200tfhe_params *p, const void *d_std, b200tfhe_bsk **out) { return bsk_convert_dev(ctx, p, 0, d_std, out); }
int b200tfhe_bsk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *s, b200tfhe_bsk **out) { return bsk_upload(ctx, p, 0, s, 64, out); }
int b200tfhe_bsk_upload_u32(b200tfhe_context *ctx, const b200tfhe_params *p, const uint32_t *s, b200tfhe_bsk **out) { return bsk_upload(ctx, p, 0, s, 32, out); }

// multi-bit keys (ShortintBootstrappingKey::MultiBit, shortint/server_key/mod.rs:103-111): same handle type, so every
// pbs / ks_pbs / blind_rotate / wave entry point accepts them and runs n/G iterations
static int check_grouping(b200tfhe_context *ctx, uint32_t grouping) {
    REQUIRE(grouping >= 2 && grouping <= 4, "multibit: grouping factor must be 2, 3 or 4");
    return 0;
}
int b200tfhe_multibit_bsk_alloc(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping, b200tfhe_bsk **out) {
    if (out) *out = nullptr;
    if (check_grouping(ctx, grouping)) return 1;
    return bsk_alloc(ctx, p, (int)grouping, out);
}
int b200tfhe_multibit_bsk_convert_dev(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping, const void *d_std, b200tfhe_bsk **out) {
    if (out) *out = nullptr;
    if (check_grouping(ctx, grouping)) return 1;
    return bsk_convert_dev(ctx, p, (int)grouping, d_std, out);
}
int b200tfhe_multibit_bsk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, uint32_t grouping, const uint64_t *s, b200tfhe_bsk **out) {
    if (out) *out = nullptr;
    if (check_grouping(ctx, grouping)) return 1;
    return bsk_upload(ctx, p, (int)grouping, s, 64, out);
}
int b200tfhe_multibit_params_supported(const b200tfhe_params *p, uint32_t grouping) {
    if (!p || grouping < 2 || grouping > 4) return 0;
    return br_mb_supported(to_params(p), (int)grouping) != 0;
}
// Fourier-domain key in the reference's serialised (natural frequency) order: [ggsw][level][row][col][N/2] c64, values as
// forward_as_torus produces them (fft/mod.rs:197-218).  For the team FFTs (N <= 2048) the device layout IS that order
// (position index == frequency index); the N = 8192 kernel keeps frequency t_hi + 16 t_lo + 256 p at position 256 p + t.
int b200tfhe_bsk_upload_fourier(b200tfhe_context *ctx, const b200tfhe_params *pp, uint32_t grouping, const double *fourier, b200tfhe_bsk **out) {
    REQUIRE(ctx && pp && out && fourier, "bsk_upload_fourier: null argument");
    *out = nullptr;
    REQUIRE(grouping >= 1 && grouping <= 4, "bsk_upload_fourier: grouping factor must be 1 (classic) .. 4");
    if (bsk_alloc(ctx, pp, grouping > 1 ? (int)grouping : 0, out)) return 1;
    b200tfhe_bsk *k = *out;
    const int R = bsk_radix(k->p, k->grouping);
    const void *src = fourier;
    std::vector<double> tmp;
    if (R == 256) {
        const size_t M = (size_t)k->p.N / 2, npoly = k->bytes / (M * sizeof(cplx));
        tmp.resize(npoly * M * 2);
        for (size_t q = 0; q < npoly; q++)
            for (size_t pos = 0; pos < M; pos++) {
                const size_t t = pos % 256, p = pos / 256, f = t / 16 + 16 * (t % 16) + 256 * p;
                tmp[2 * (q * M + pos)] = fourier[2 * (q * M + f)];
                tmp[2 * (q * M + pos) + 1] = fourier[2 * (q * M + f) + 1];
            }
        src = tmp.data();
    }
    cudaError_t e = cudaMemcpy(k->d_fourier, src, k->bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { b200tfhe_bsk_destroy(k); *out = nullptr; return fail_cuda(ctx, "cudaMemcpy(fourier bsk)", e); }
    return 0;
}
uint32_t b200tfhe_bsk_grouping_factor(const b200tfhe_bsk *bsk) { return bsk && bsk->grouping ? (uint32_t)bsk->grouping : 1u; }

int b200tfhe_bsk_device_buffer(const b200tfhe_bsk *bsk, void **d_ptr, size_t *bytes) {
    if (!bsk || !d_ptr || !bytes) return 1;
    *d_ptr = bsk->d_fourier;
    *bytes = bsk->bytes;
    return 0;
}
int b200tfhe_bsk_destroy(b200tfhe_bsk *bsk) {
    if (!bsk) return 0;
    cudaSetDevice(bsk->device);
    cudaFree(bsk->d_fourier);
    delete bsk;
    return 0;
}

int b200tfhe_ksk_alloc(b200tfhe_context *ctx, const b200tfhe_params *pp, b200tfhe_ksk **out) {
    REQUIRE(ctx && pp && out, "ksk_alloc: null argument");
    *out = nullptr;
    Params p = to_params(pp);
    if (check_ks_params(ctx, p)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    size_t bytes = (size_t)p.k * p.N * p.ks_level * small_len(p) * sz(p);
    b200tfhe_ksk *k = new b200tfhe_ksk{ctx, p, nullptr, nullptr, bytes, false, {}};
    k->device = ctx->device;
    cudaError_t e = cudaMalloc(&k->d_ksk, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&k->d_colsum, small_len(p) * sz(p));
    if (e != cudaSuccess) { cudaFree(k->d_ksk); delete k; return fail_cuda(ctx, "cudaMalloc(ksk)", e); }
    *out = k;
    return 0;
}
int b200tfhe_ksk_finalize(b200tfhe_context *ctx, b200tfhe_ksk *k) {
    REQUIRE(ctx && k, "ksk_finalize: null argument");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const int n_in = k->p.k * k->p.N;
    int rc;
    if (k->p.ks_base_log > 7) {
        // wide base (e.g. 2^10 / 2^11 of the 3_3 / 4_4 WoP-PBS sets): key rows replicated with shifts for the exact digit split
        const int bl = k->p.ks_base_log, sub = (bl + 5) / 6, sub_bits = (bl + sub - 1) / sub, rows_per = k->p.ks_level * sub;   // sub-digits of <= 6 bits (+ carry) fit int8
        REQUIRE(ks_tc_supported(k->p.bits, n_in, k->p.n, rows_per, sub_bits + 1), "keyswitch base_log > 7 needs the tensor-core path, which does not cover this set");
        const size_t width = (size_t)k->p.n + 1;
        void *d_exp = nullptr, *planes = nullptr;
        CK(cudaMalloc(&d_exp, (size_t)n_in * rows_per * width * sz(k->p)));
        cudaError_t e = cudaMalloc(&planes, ks_tc_key_bytes(k->p.bits, n_in, k->p.n, rows_per));
        if (e != cudaSuccess) { cudaFree(d_exp); return fail_cuda(ctx, "cudaMalloc(ksk planes)", e); }
        if (k->tk.d_planes) { cudaFree(k->tk.d_planes); k->tk.d_planes = nullptr; }
        rc = ks_expand_rows_launch(k->d_ksk, d_exp, (size_t)n_in, k->p.ks_level, sub, sub_bits, 0, width, k->p.bits, ctx->stream);
        if (!rc) rc = ks_tc_prepare_key(&k->tk, d_exp, planes, k->p.bits, n_in, k->p.n, rows_per, ctx->stream);
        ctx->launches += 2;
        cudaError_t es = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_exp);
        if (rc || es != cudaSuccess) { cudaFree(planes); k->tk.d_planes = nullptr; return fail(ctx, "wide keyswitch key preparation failed: " + std::to_string(rc)); }
        k->sub = sub;
        k->sub_bits = sub_bits;
        k->finalized = true;
        return 0;
    }
    rc = ks_colsum_launch(k->d_ksk, k->d_colsum, k->p.k * k->p.N, k->p.n, k->p.ks_level, k->p.ks_base_log, k->p.bits, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "ks_colsum launch failed: " + std::to_string(rc));
    if (ks_tc_supported(k->p.bits, n_in, k->p.n, k->p.ks_level, k->p.ks_base_log)) {
        if (k->tk.d_planes) { cudaFree(k->tk.d_planes); k->tk.d_planes = nullptr; }
        void *planes = nullptr;
        CK(cudaMalloc(&planes, ks_tc_key_bytes(k->p.bits, n_in, k->p.n, k->p.ks_level)));
        rc = ks_tc_prepare_key(&k->tk, k->d_ksk, planes, k->p.bits, n_in, k->p.n, k->p.ks_level, ctx->stream);
        ctx->launches++;
        if (rc) { cudaFree(planes); k->tk.d_planes = nullptr; return fail(ctx, "ks_tc_prepare_key failed: " + std::to_string(rc)); }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    k->finalized = true;
    return 0;
}
static int ksk_upload(b200tfhe_context *ctx, const b200tfhe_params *pp, const void *ksk, int bits, b200tfhe_ksk **out) {
    REQUIRE(ctx && pp && out && ksk, "ksk_upload: null argument");
    *out = nullptr;
    REQUIRE(pp->torus_bits == bits, "ksk_upload: torus_bits does not match the entry point");
    if (b200tfhe_ksk_alloc(ctx, pp, out)) return 1;
    cudaError_t e = cudaMemcpy((*out)->d_ksk, ksk, (*out)->bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { b200tfhe_ksk_destroy(*out); *out = nullptr; return fail_cuda(ctx, "cudaMemcpy(ksk)", e); }
    if (b200tfhe_ksk_finalize(ctx, *out)) { b200tfhe_ksk_destroy(*out); *out = nullptr; return 1; }
    return 0;
}
int b200tfhe_ksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *p, const uint64_t *k, b200tfhe_ksk **out) { return ksk_upload(ctx, p, k, 64, out); }
int b200tfhe_ksk_upload_u32(b200tfhe_context *ctx, const b200tfhe_params *p, const uint32_t *k, b200tfhe_ksk **out) { return ksk_upload(ctx, p, k, 32, out); }
int b200tfhe_ksk_device_buffer(const b200tfhe_ksk *ksk, void **d_ptr, size_t *bytes) {
    if (!ksk || !d_ptr || !bytes) return 1;
    *d_ptr = ksk->d_ksk;
    *bytes = ksk->bytes;
    return 0;
}
int b200tfhe_ksk_destroy(b200tfhe_ksk *ksk) {
    if (!ksk) return 0;
    cudaSetDevice(ksk->device);
    cudaFree(ksk->d_ksk);
    cudaFree(ksk->d_colsum);
    if (ksk->tk.d_planes) cudaFree(ksk->tk.d_planes);
    delete ksk;
    return 0;
}

}  // extern "C"

// ---- device-pointer operations (caller holds ctx->mu) ---------------------------------------------------------
static int ks_dev_locked(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const void *d_in, void *d_out, uint32_t batch) {
    REQUIRE(ksk->finalized, "keyswitch: key not finalized (call b200tfhe_ksk_finalize after filling the device buffer)");
    const Params &p = ksk->p;
    KSLaunch a{d_in, big_len(p), d_out, small_len(p), ksk->d_ksk, ksk->d_colsum, p.k * p.N, p.n, p.ks_level, p.ks_base_log, p.bits, (int)batch};
    a.body_col = ksk->body_col;
    a.sub = ksk->sub;
    a.sub_bits = ksk->sub_bits;
    int rc;
    if (ksk->tk.d_planes && (((int)batch >= ctx->opt_ks_tc_min && ctx->opt_ks_tensor_core) || ksk->sub > 1)) {
        void *digits = nullptr;
        CK(ctx->ws.get(Workspace::DIG, (size_t)batch * ksk->tk.kp, &digits));
        rc = ks_tc_launch(a, ksk->tk, digits, ctx->stream);
        ctx->launches += 2;
    } else {
        rc = ks_launch(a, ctx->stream);
        ctx->launches += a.launches;
    }
    if (rc) return fail(ctx, "keyswitch launch failed: " + std::to_string(rc));
    return 0;
}
static int br_dev_locked(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const void *d_in, void *d_out, const void *d_luts,
                         const uint32_t *d_idx, uint32_t batch, int mode) {
    const Params &p = bsk->p;
    int R = bsk_radix(p, bsk->grouping);
    BRLaunch a{d_in, small_len(p), d_luts, d_idx, bsk->d_fourier, d_out, mode == BR_MODE_GLWE ? glwe_len(p) : big_len(p),
               p.n, (int)batch, mode, tw_fwd(ctx, R), tw_inv(ctx, R)};
    a.force = ctx->opt_br_kernel;
    a.lat_max = ctx->opt_lat_max;
    a.lat_cluster_max = ctx->opt_lat_cluster_max;
    if (bsk->grouping && (int)batch <= ctx->opt_mb_lat_max) {
        if (const size_t need = br_mb_lat_scratch_bytes(p, bsk->grouping, (int)batch)) {
            CK(ctx->ws.get(Workspace::MBK, need, &a.mb_scratch));
            a.mb_lat_max = ctx->opt_mb_lat_max;
        }
    }
    int rc = bsk->grouping ? br_mb_launch(p, bsk->grouping, a, ctx->stream) : br_launch(p, a, ctx->stream);
    ctx->launches += a.launches;
    if (rc == -2) return fail(ctx, "blind_rotate: the kernel selected with option \"br_kernel\" is not built for this parameter set");
    if (rc) return fail(ctx, "blind_rotate launch failed: " + std::to_string(rc));
    return 0;
}
static int check_pair(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk) {
    REQUIRE(ksk && bsk, "null key handle");
    REQUIRE(ksk->ctx == ctx && bsk->ctx == ctx, "key handle belongs to another context");
    const Params &a = ksk->p, &b = bsk->p;
    // the reference asserts matching dimensions / moduli: lwe_programmable_bootstrapping.rs:1083-1097
    REQUIRE(a.n == b.n && a.k == b.k && a.N == b.N && a.bits == b.bits, "keyswitch key and bootstrap key dimensions differ");
    return 0;
}

extern "C" {

int b200tfhe_keyswitch_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const void *d_in, void *d_out, uint32_t batch) {
    REQUIRE(ctx && ksk && (batch == 0 || (d_in && d_out)), "keyswitch_batch_dev: null argument");
    REQUIRE(ksk->ctx == ctx, "key handle belongs to another context");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return ks_dev_locked(ctx, ksk, d_in, d_out, batch);
}
// Allocation-free variant of keyswitch_batch_dev (no counterpart in the reference; see noalloc.cu): the digit scratch of the
// tensor-core keyswitch is a fixed __device__ array instead of the context's growable workspace.
static_assert(NOALLOC_KS_SCRATCH_BYTES == B200TFHE_NOALLOC_KS_SCRATCH_BYTES, "scratch size: kernels.h vs b200tfhe.h");
int b200tfhe_keyswitch_batch_noalloc_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const void *d_in, void *d_out, uint32_t batch) {
    REQUIRE(ctx && ksk && (batch == 0 || (d_in && d_out)), "keyswitch_batch_noalloc_dev: null argument");
    REQUIRE(ksk->ctx == ctx, "key handle belongs to another context");
    REQUIRE(ksk->finalized, "keyswitch: key not finalized (call b200tfhe_ksk_finalize after filling the device buffer)");
    REQUIRE(ksk->tk.d_planes, "keyswitch_batch_noalloc_dev: this key has no tensor-core byte planes; b200tfhe_keyswitch_batch_dev "
                              "runs the CUDA-core kernel for it, which needs no scratch");
    const Params &p = ksk->p;
    const size_t max_batch = NOALLOC_KS_SCRATCH_BYTES / ksk->tk.kp;
    REQUIRE((size_t)batch <= max_batch,
            "keyswitch_batch_noalloc_dev: batch " + std::to_string(batch) + " needs " + std::to_string((size_t)batch * ksk->tk.kp) +
                " bytes of digit scratch, the fixed device array holds " + std::to_string((size_t)NOALLOC_KS_SCRATCH_BYTES) +
                " (at most " + std::to_string(max_batch) + " ciphertexts per call for this key)");
    if (batch == 0) return 0;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    KSLaunch a{d_in, big_len(p), d_out, small_len(p), ksk->d_ksk, ksk->d_colsum, p.k * p.N, p.n, p.ks_level, p.ks_base_log, p.bits, (int)batch};
    a.body_col = ksk->body_col;
    a.sub = ksk->sub;
    a.sub_bits = ksk->sub_bits;
    cudaError_t ce = cudaSuccess;
    const int rc = ks_tc_noalloc_launch(a, ksk->tk, ctx->device, ctx->stream, &ce);
    if (rc == -12) return fail_cuda(ctx, "cudaGetSymbolAddress(noalloc keyswitch scratch)", ce);
    if (rc) return fail(ctx, "keyswitch (noalloc) launch failed: " + std::to_string(rc));
    ctx->launches += 2;
    return 0;
}
int b200tfhe_pbs_batch_dev(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const void *d_in, void *d_out, const void *d_luts,
                           const uint32_t *d_idx, uint32_t batch) {
    REQUIRE(ctx && bsk && (batch == 0 || (d_in && d_out && d_luts)), "pbs_batch_dev: null argument");
    REQUIRE(bsk->ctx == ctx, "key handle belongs to another context");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return br_dev_locked(ctx, bsk, d_in, d_out, d_luts, d_idx, batch, BR_MODE_LWE);
}
int b200tfhe_blind_rotate_batch_dev(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const void *d_in, void *d_out, const void *d_luts,
                                    const uint32_t *d_idx, uint32_t batch) {
    REQUIRE(ctx && bsk && (batch == 0 || (d_in && d_out && d_luts)), "blind_rotate_batch_dev: null argument");
    REQUIRE(bsk->ctx == ctx, "key handle belongs to another context");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return br_dev_locked(ctx, bsk, d_in, d_out, d_luts, d_idx, batch, BR_MODE_GLWE);
}
int b200tfhe_ks_pbs_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const void *d_in, void *d_out,
                              const void *d_luts, const uint32_t *d_idx, uint32_t batch) {
    REQUIRE(ctx && (batch == 0 || (d_in && d_out && d_luts)), "ks_pbs_batch_dev: null argument");
    if (check_pair(ctx, ksk, bsk)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *d_mid;
    CK(ctx->ws.get(Workspace::MID, (size_t)batch * small_len(bsk->p) * sz(bsk->p), &d_mid));
    if (ks_dev_locked(ctx, ksk, d_in, d_mid, batch)) return 1;
    return br_dev_locked(ctx, bsk, d_mid, d_out, d_luts, d_idx, batch, BR_MODE_LWE);
}
int b200tfhe_pbs_ks_batch_dev(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const void *d_in, void *d_out,
                              const void *d_luts, const uint32_t *d_idx, uint32_t batch) {
    REQUIRE(ctx && (batch == 0 || (d_in && d_out && d_luts)), "pbs_ks_batch_dev: null argument");
    if (check_pair(ctx, ksk, bsk)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *d_mid;
    CK(ctx->ws.get(Workspace::MID, (size_t)batch * big_len(bsk->p) * sz(bsk->p), &d_mid));
    if (br_dev_locked(ctx, bsk, d_in, d_mid, d_luts, d_idx, batch, BR_MODE_LWE)) return 1;
    return ks_dev_locked(ctx, ksk, d_mid, d_out, batch);
}

}  // extern "C"

// ---- host-pointer operations ---------------------------------------------------------------------------------
enum Op { OP_KS, OP_PBS, OP_BR, OP_KS_PBS, OP_PBS_KS };

static int host_op(b200tfhe_context *ctx, Op op, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, int bits, const void *in,
                   void *out, const void *luts, uint32_t n_luts, const uint32_t *lut_idx, uint32_t batch) {
    REQUIRE(ctx, "null context");
    const bool need_ksk = (op == OP_KS || op == OP_KS_PBS || op == OP_PBS_KS), need_bsk = (op != OP_KS);
    REQUIRE(!need_ksk || ksk, "null keyswitch key");
    REQUIRE(!need_bsk || bsk, "null bootstrap key");
    if (need_ksk && need_bsk && check_pair(ctx, ksk, bsk)) return 1;
    REQUIRE(!need_ksk || ksk->ctx == ctx, "key handle belongs to another context");
    REQUIRE(!need_bsk || bsk->ctx == ctx, "key handle belongs to another context");
    const Params &p = need_bsk ? bsk->p : ksk->p;
    REQUIRE(p.bits == bits, "torus width of the key does not match the entry point");
    if (batch == 0) return 0;
    REQUIRE(in && out, "null ciphertext buffer");
    REQUIRE(!need_bsk || (luts && n_luts > 0), "null / empty accumulator list");
    if (need_bsk && lut_idx)
        for (uint32_t i = 0; i < batch; i++) REQUIRE(lut_idx[i] < n_luts, "lut_idx out of range");

    const size_t s = sz(p);
    const size_t in_len = (op == OP_KS || op == OP_KS_PBS) ? big_len(p) : small_len(p);
    const size_t out_len = (op == OP_KS || op == OP_PBS_KS) ? small_len(p) : (op == OP_BR ? glwe_len(p) : big_len(p));
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    DrainOnExit drain(ctx);
    void *d_in, *d_out, *d_mid = nullptr, *d_luts = nullptr, *d_idx = nullptr;
    CK(ctx->ws.get(Workspace::IN, batch * in_len * s, &d_in));
    CK(ctx->ws.get(Workspace::OUT, batch * out_len * s, &d_out));
    if (need_bsk) {
        CK(ctx->ws.get(Workspace::LUTS, (size_t)n_luts * glwe_len(p) * s, &d_luts));
        CK(cudaMemcpyAsync(d_luts, luts, (size_t)n_luts * glwe_len(p) * s, cudaMemcpyHostToDevice, ctx->stream));
        if (lut_idx) {
            CK(ctx->ws.get(Workspace::IDX, (size_t)batch * 4, &d_idx));
            CK(cudaMemcpyAsync(d_idx, lut_idx, (size_t)batch * 4, cudaMemcpyHostToDevice, ctx->stream));
        }
    }
    // one chunk of the batch through the kernels (device pointers already offset)
    auto run = [&](const char *ci, char *co, const uint32_t *cidx, uint32_t cb, size_t first) -> int {
        int rc = 0;
        switch (op) {
            case OP_KS: rc = ks_dev_locked(ctx, ksk, ci, co, cb); break;
            case OP_PBS: rc = br_dev_locked(ctx, bsk, ci, co, d_luts, cidx, cb, BR_MODE_LWE); break;
            case OP_BR: rc = br_dev_locked(ctx, bsk, ci, co, d_luts, cidx, cb, BR_MODE_GLWE); break;
            case OP_KS_PBS: {
                char *mid = (char *)d_mid + first * small_len(p) * s;
                rc = ks_dev_locked(ctx, ksk, ci, mid, cb);
                if (!rc) rc = br_dev_locked(ctx, bsk, mid, co, d_luts, cidx, cb, BR_MODE_LWE);
                break;
            }
            case OP_PBS_KS: {
                char *mid = (char *)d_mid + first * big_len(p) * s;
                rc = br_dev_locked(ctx, bsk, ci, mid, d_luts, cidx, cb, BR_MODE_LWE);
                if (!rc) rc = ks_dev_locked(ctx, ksk, mid, co, cb);
                break;
            }
        }
        return rc;
    };
    if (op == OP_KS_PBS) CK(ctx->ws.get(Workspace::MID, batch * small_len(p) * s, &d_mid));
    if (op == OP_PBS_KS) CK(ctx->ws.get(Workspace::MID, batch * big_len(p) * s, &d_mid));

    // chunking: H2D of chunk i+1 and D2H of chunk i-1 overlap the kernels of chunk i.  Chunks are whole waves of the blind-rotation
    // kernel that will run (unit), the FIRST and the LAST chunk are one wave each -- what stays exposed is the copy-in of the first
    // chunk and the copy-out of the last, so they are the smallest -- and the middle is cut into at most three pieces.
    const uint32_t min_chunk = ctx->opt_chunk_min;
    uint32_t bounds[b200tfhe_context::MAX_CHUNKS + 1] = {0};
    uint32_t n_chunks = 1;
    bounds[1] = batch;
    if (batch >= 2 * min_chunk && min_chunk > 0) {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        const uint32_t unit = need_bsk ? (uint32_t)((br_tmem_supported(p) ? 4 : br_x16_supported(p) ? 8 : br_big_supported(p) ? 1 : 4) * sms) : 592u;
        const uint32_t units = batch / unit, rem = batch % unit;
        if (units >= 6) {
            const uint32_t last = rem >= unit / 4 ? rem : unit + rem;           // a very small last chunk would waste a launch
            const uint32_t mid_units = (batch - unit - last) / unit;
            const uint32_t pieces = mid_units >= 3 ? 3 : (mid_units >= 1 ? mid_units : 0);
            n_chunks = 0;
            bounds[++n_chunks] = unit;
            for (uint32_t q = 0; q < pieces; q++) bounds[n_chunks + 1] = bounds[n_chunks] + (mid_units / pieces + (q < mid_units % pieces ? 1 : 0)) * unit, n_chunks++;
            bounds[++n_chunks] = batch;
        } else if (units >= 2) {
            n_chunks = 2;
            bounds[1] = (units / 2) * unit;
            bounds[2] = batch;
        }
    }
    if (n_chunks == 1) {
        CK(cudaMemcpyAsync(d_in, in, batch * in_len * s, cudaMemcpyHostToDevice, ctx->stream));
        if (run((const char *)d_in, (char *)d_out, (const uint32_t *)d_idx, batch, 0)) return 1;
        CK(cudaMemcpyAsync(out, d_out, batch * out_len * s, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        drain.armed = false;
        return 0;
    }
    if (!ctx->s_h2d) {
        CK(cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->ev_start, cudaEventDisableTiming));
        for (int i = 0; i < b200tfhe_context::MAX_CHUNKS; i++) {
            CK(cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
        }
    }
    // the copy streams start after whatever is already queued on the main stream (LUT / index uploads, earlier calls)
    CK(cudaEventRecord(ctx->ev_start, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->s_h2d, ctx->ev_start, 0));
    CK(cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_start, 0));
    for (uint32_t c = 0; c < n_chunks; c++) {
        const size_t first = bounds[c];
        const uint32_t cb = bounds[c + 1] - bounds[c];
        CK(cudaMemcpyAsync((char *)d_in + first * in_len * s, (const char *)in + first * in_len * s, (size_t)cb * in_len * s,
                           cudaMemcpyHostToDevice, ctx->s_h2d));
        CK(cudaEventRecord(ctx->ev_in[c], ctx->s_h2d));
    }
    for (uint32_t c = 0; c < n_chunks; c++) {
        const size_t first = bounds[c];
        const uint32_t cb = bounds[c + 1] - bounds[c];
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_in[c], 0));
        if (run((const char *)d_in + first * in_len * s, (char *)d_out + first * out_len * s,
                d_idx ? (const uint32_t *)d_idx + first : nullptr, cb, first))
            return 1;
        CK(cudaEventRecord(ctx->ev_done[c], ctx->stream));
        CK(cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_done[c], 0));
        CK(cudaMemcpyAsync((char *)out + first * out_len * s, (const char *)d_out + first * out_len * s, (size_t)cb * out_len * s,
                           cudaMemcpyDeviceToHost, ctx->s_d2h));
    }
    CK(cudaStreamSynchronize(ctx->s_d2h));
    CK(cudaStreamSynchronize(ctx->stream));
    drain.armed = false;
    return 0;
}

extern "C" {

#define HOST_OPS(SUF, T, BITS)                                                                                                          \
    int b200tfhe_keyswitch_batch_##SUF(b200tfhe_context *c, const b200tfhe_ksk *k, const T *in, T *out, uint32_t b) {                   \
        return host_op(c, OP_KS, k, nullptr, BITS, in, out, nullptr, 0, nullptr, b);                                                    \
    }                                                                                                                                   \
    int b200tfhe_pbs_batch_##SUF(b200tfhe_context *c, const b200tfhe_bsk *k, const T *in, T *out, const T *l, uint32_t nl,              \
                                 const uint32_t *li, uint32_t b) {                                                                      \
        return host_op(c, OP_PBS, nullptr, k, BITS, in, out, l, nl, li, b);                                                             \
    }                                                                                                                                   \
    int b200tfhe_blind_rotate_batch_##SUF(b200tfhe_context *c, const b200tfhe_bsk *k, const T *in, T *out, const T *l, uint32_t nl,     \
                                          const uint32_t *li, uint32_t b) {                                                             \
        return host_op(c, OP_BR, nullptr, k, BITS, in, out, l, nl, li, b);                                                              \
    }                                                                                                                                   \
    int b200tfhe_ks_pbs_batch_##SUF(b200tfhe_context *c, const b200tfhe_ksk *ks, const b200tfhe_bsk *k, const T *in, T *out,            \
                                    const T *l, uint32_t nl, const uint32_t *li, uint32_t b) {                                          \
        return host_op(c, OP_KS_PBS, ks, k, BITS, in, out, l, nl, li, b);                                                               \
    }                                                                                                                                   \
    int b200tfhe_pbs_ks_batch_##SUF(b200tfhe_context *c, const b200tfhe_ksk *ks, const b200tfhe_bsk *k, const T *in, T *out,            \
                                    const T *l, uint32_t nl, const uint32_t *li, uint32_t b) {                                          \
        return host_op(c, OP_PBS_KS, ks, k, BITS, in, out, l, nl, li, b);                                                               \
    }
HOST_OPS(u64, uint64_t, 64)
HOST_OPS(u32, uint32_t, 32)
#undef HOST_OPS

// add_external_product_assign / cmux_assign over a list (a10): host buffers staged through the workspace, one CTA per ciphertext
static int external_product_host(b200tfhe_context *ctx, const b200tfhe_bsk *ggsw, const uint32_t *idx, void *out, void *src, uint32_t batch,
                                 int cmux, int bits) {
    REQUIRE(ctx && ggsw, "external_product: null argument");
    REQUIRE(ggsw->ctx == ctx, "key handle belongs to another context");
    REQUIRE(ggsw->grouping == 0, "external_product: a multi-bit bootstrap key is not a GGSW list");
    const Params &p = ggsw->p;
    REQUIRE(p.bits == bits, "torus width of the GGSW list does not match the entry point");
    const int R = external_product_supported(p);
    REQUIRE(R != 0, "external product is not built for this (N, k, level, base_log, bits) set");
    if (batch == 0) return 0;
    REQUIRE(out && src, "null ciphertext buffer");
    if (idx)
        for (uint32_t i = 0; i < batch; i++) REQUIRE(idx[i] < (uint32_t)p.n, "ggsw_idx out of range");
    const size_t bytes = (size_t)batch * glwe_len(p) * sz(p);
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *d_out, *d_src, *d_idx = nullptr;
    CK(ctx->ws.get(Workspace::OUT, bytes, &d_out));
    CK(ctx->ws.get(Workspace::IN, bytes, &d_src));
    CK(cudaMemcpyAsync(d_out, out, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_src, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (idx) {
        CK(ctx->ws.get(Workspace::IDX, (size_t)batch * 4, &d_idx));
        CK(cudaMemcpyAsync(d_idx, idx, (size_t)batch * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    const int rc = external_product_launch(p, d_out, d_src, ggsw->d_fourier, (const uint32_t *)d_idx, cmux, (int)batch, tw_fwd(ctx, R),
                                           tw_inv(ctx, R), ctx->stream);
    ctx->launches++;
    if (rc) {
        cudaStreamSynchronize(ctx->stream);
        return fail(ctx, "external_product launch failed: " + std::to_string(rc));
    }
    cudaError_t e = cudaMemcpyAsync(out, d_out, bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && cmux) e = cudaMemcpyAsync(src, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream);
    cudaError_t es = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess || es != cudaSuccess) return fail_cuda(ctx, "external_product", e != cudaSuccess ? e : es);
    return 0;
}
int b200tfhe_external_product_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *g, const uint32_t *idx, uint64_t *out, const uint64_t *glwe,
                                        uint32_t batch) {
    return external_product_host(ctx, g, idx, out, const_cast<uint64_t *>(glwe), batch, 0, 64);
}
int b200tfhe_external_product_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *g, const uint32_t *idx, uint32_t *out, const uint32_t *glwe,
                                        uint32_t batch) {
    return external_product_host(ctx, g, idx, out, const_cast<uint32_t *>(glwe), batch, 0, 32);
}
int b200tfhe_cmux_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *g, const uint32_t *idx, uint64_t *ct0, uint64_t *ct1, uint32_t batch) {
    return external_product_host(ctx, g, idx, ct0, ct1, batch, 1, 64);
}
int b200tfhe_cmux_batch_u32(b200tfhe_context *ctx, const b200tfhe_bsk *g, const uint32_t *idx, uint32_t *ct0, uint32_t *ct1, uint32_t batch) {
    return external_product_host(ctx, g, idx, ct0, ct1, batch, 1, 32);
}

int b200tfhe_gate_batch_u32(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, int gate, int pbs_first,
                            const uint32_t *a, const uint32_t *b, uint32_t *out, uint32_t batch) {
    REQUIRE(ctx, "null context");
    if (check_pair(ctx, ksk, bsk)) return 1;
    REQUIRE(gate >= 0 && gate <= 5, "unknown gate");
    const Params &p = bsk->p;
    REQUIRE(p.bits == 32, "boolean gates are defined on the 2^32 torus");
    if (batch == 0) return 0;
    REQUIRE(a && b && out, "null ciphertext buffer");
    const size_t len = pbs_first ? small_len(p) : big_len(p);
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *d_a, *d_b, *d_in, *d_mid, *d_out, *d_lut;
    CK(ctx->ws.get(Workspace::A, batch * len * 4, &d_a));
    CK(ctx->ws.get(Workspace::B, batch * len * 4, &d_b));
    CK(ctx->ws.get(Workspace::IN, batch * len * 4, &d_in));
    CK(ctx->ws.get(Workspace::OUT, batch * len * 4, &d_out));
    CK(ctx->ws.get(Workspace::MID, batch * (pbs_first ? big_len(p) : small_len(p)) * 4, &d_mid));
    CK(ctx->ws.get(Workspace::LUTS, glwe_len(p) * 4, &d_lut));
    CK(cudaMemcpyAsync(d_a, a, batch * len * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_b, b, batch * len * 4, cudaMemcpyHostToDevice, ctx->stream));
    {
        // constant accumulator: mask 0, body all PLAINTEXT_TRUE (boolean/engine/bootstrapping.rs:59-60)
        std::vector<uint32_t> lut(glwe_len(p), 0u);
        for (int j = 0; j < p.N; j++) lut[(size_t)p.k * p.N + j] = 1u << 29;
        CK(cudaMemcpyAsync(d_lut, lut.data(), lut.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));  // lut is a stack-lifetime host buffer
    }
    int rc = gate_prologue_launch(gate, (const uint32_t *)d_a, (const uint32_t *)d_b, (uint32_t *)d_in, (int)len, (int)batch, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "gate prologue launch failed: " + std::to_string(rc));
    if (pbs_first) {
        rc = br_dev_locked(ctx, bsk, d_in, d_mid, d_lut, nullptr, batch, BR_MODE_LWE);
        if (!rc) rc = ks_dev_locked(ctx, ksk, d_mid, d_out, batch);
    } else {
        rc = ks_dev_locked(ctx, ksk, d_in, d_mid, batch);
        if (!rc) rc = br_dev_locked(ctx, bsk, d_mid, d_out, d_lut, nullptr, batch, BR_MODE_LWE);
    }
    if (rc) return 1;
    CK(cudaMemcpyAsync(out, d_out, batch * len * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

static int modswitch_host(b200tfhe_context *ctx, const void *in, uint32_t *out, size_t count, uint32_t log2N, int bits) {
    REQUIRE(ctx, "null context");
    if (count == 0) return 0;
    REQUIRE(in && out, "null buffer");
    REQUIRE(log2N >= 1 && (int)log2N + 2 <= bits, "invalid polynomial size");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *d_in, *d_out;
    CK(ctx->ws.get(Workspace::IN, count * (bits / 8), &d_in));
    CK(ctx->ws.get(Workspace::OUT, count * 4, &d_out));
    CK(cudaMemcpyAsync(d_in, in, count * (bits / 8), cudaMemcpyHostToDevice, ctx->stream));
    int rc = modswitch_launch(d_in, (uint32_t *)d_out, count, (int)log2N, bits, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "modswitch launch failed");
    CK(cudaMemcpyAsync(out, d_out, count * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int b200tfhe_modulus_switch_u64(b200tfhe_context *ctx, const uint64_t *in, uint32_t *out, size_t count, uint32_t l) { return modswitch_host(ctx, in, out, count, l, 64); }
int b200tfhe_modulus_switch_u32(b200tfhe_context *ctx, const uint32_t *in, uint32_t *out, size_t count, uint32_t l) { return modswitch_host(ctx, in, out, count, l, 32); }

static int sample_extract_host(b200tfhe_context *ctx, const void *glwe, void *lwe, uint32_t k, uint32_t N, uint32_t batch, int bits) {
    REQUIRE(ctx, "null context");
    if (batch == 0) return 0;
    REQUIRE(glwe && lwe && k > 0 && N > 0, "invalid argument");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t s = bits / 8, in_b = (size_t)batch * (k + 1) * N * s, out_b = (size_t)batch * ((size_t)k * N + 1) * s;
    void *d_in, *d_out;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, glwe, in_b, cudaMemcpyHostToDevice, ctx->stream));
    int rc = sample_extract_launch(d_in, d_out, (int)k, (int)N, (int)batch, bits, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "sample_extract launch failed");
    CK(cudaMemcpyAsync(lwe, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int b200tfhe_sample_extract_batch_u64(b200tfhe_context *ctx, const uint64_t *g, uint64_t *l, uint32_t k, uint32_t N, uint32_t b) { return sample_extract_host(ctx, g, l, k, N, b, 64); }
int b200tfhe_sample_extract_batch_u32(b200tfhe_context *ctx, const uint32_t *g, uint32_t *l, uint32_t k, uint32_t N, uint32_t b) { return sample_extract_host(ctx, g, l, k, N, b, 32); }



// ---- gadget extras: packing keyswitch, multi-value bootstrap, accumulator packing -------------------
int b200tfhe_pksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *pp, const uint64_t *pksk, b200tfhe_pksk **out) {
    REQUIRE(ctx && pp && out && pksk, "pksk_upload: null argument");
    *out = nullptr;
    REQUIRE(pp->torus_bits == 64, "pksk_upload: 64-bit torus only");
    REQUIRE(pp->glwe_dimension > 0 && is_pow2(pp->polynomial_size), "invalid dimensions");
    b200tfhe_params q = *pp;
    q.lwe_dimension = (pp->glwe_dimension + 1) * pp->polynomial_size - 1;    // output "LWE" = the flat GLWE
    b200tfhe_ksk *inner = nullptr;
    if (b200tfhe_ksk_upload_u64(ctx, &q, pksk, &inner)) return 1;
    inner->body_col = pp->glwe_dimension * pp->polynomial_size;
    *out = new b200tfhe_pksk{inner, to_params(pp)};
    return 0;
}
int b200tfhe_pksk_destroy(b200tfhe_pksk *k) {
    if (!k) return 0;
    b200tfhe_ksk_destroy(k->inner);
    delete k;
    return 0;
}

int b200tfhe_packing_keyswitch_batch_u64(b200tfhe_context *ctx, const b200tfhe_pksk *pksk, const uint64_t *in, uint64_t *out_glwe, uint32_t batch) {
    REQUIRE(ctx && pksk, "packing_keyswitch: null argument");
    REQUIRE(pksk->inner->ctx == ctx, "key handle belongs to another context");
    if (batch == 0) return 0;
    REQUIRE(in && out_glwe, "packing_keyswitch: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const Params &p = pksk->p;
    const size_t in_b = (size_t)batch * big_len(p) * 8, out_b = (size_t)batch * glwe_len(p) * 8;
    void *d_in, *d_out;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, in, in_b, cudaMemcpyHostToDevice, ctx->stream));
    if (ks_dev_locked(ctx, pksk->inner, d_in, d_out, batch)) return 1;
    CK(cudaMemcpyAsync(out_glwe, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int b200tfhe_glwe_mul_plain_extract_batch_u64(b200tfhe_context *ctx, const b200tfhe_params *pp, const uint64_t *glwe, const uint64_t *polys,
                                              uint32_t n_polys, uint64_t *out, uint32_t batch) {
    REQUIRE(ctx && pp, "glwe_mul_plain_extract: null argument");
    if (batch == 0 || n_polys == 0) return 0;
    REQUIRE(glwe && polys && out, "glwe_mul_plain_extract: null buffer");
    REQUIRE(pp->glwe_dimension > 0 && is_pow2(pp->polynomial_size) && pp->polynomial_size <= 8192, "invalid dimensions");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    Params p = to_params(pp);
    const size_t in_b = (size_t)batch * glwe_len(p) * 8, pl_b = (size_t)n_polys * p.N * 8, out_b = (size_t)batch * n_polys * big_len(p) * 8;
    void *d_in, *d_pl, *d_out;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::A, pl_b, &d_pl));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, glwe, in_b, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_pl, polys, pl_b, cudaMemcpyHostToDevice, ctx->stream));
    int rc = glwe_mul_plain_extract_launch(d_in, d_pl, d_out, p.k, p.N, (int)n_polys, (int)batch, 64, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "glwe_mul_plain_extract launch failed: " + std::to_string(rc));
    CK(cudaMemcpyAsync(out, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int b200tfhe_mvb_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                           const uint64_t *common_acc, const uint64_t *polys, uint32_t n_polys, uint64_t *out, uint32_t batch) {
    REQUIRE(ctx && bsk, "mvb: null argument");
    REQUIRE(bsk->ctx == ctx, "key handle belongs to another context");
    REQUIRE(bsk->p.bits == 64, "mvb: 64-bit torus only");
    if (ksk && check_pair(ctx, ksk, bsk)) return 1;
    if (batch == 0 || n_polys == 0) return 0;
    REQUIRE(in && common_acc && polys && out, "mvb: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const Params &p = bsk->p;
    const size_t in_len = ksk ? big_len(p) : small_len(p);
    const size_t in_b = (size_t)batch * in_len * 8, pl_b = (size_t)n_polys * p.N * 8, out_b = (size_t)batch * n_polys * big_len(p) * 8;
    void *d_in, *d_lut, *d_glwe, *d_pl, *d_out, *d_small = nullptr;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::LUTS, glwe_len(p) * 8, &d_lut));
    CK(ctx->ws.get(Workspace::MID, (size_t)batch * glwe_len(p) * 8, &d_glwe));
    CK(ctx->ws.get(Workspace::A, pl_b, &d_pl));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, in, in_b, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_lut, common_acc, glwe_len(p) * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_pl, polys, pl_b, cudaMemcpyHostToDevice, ctx->stream));
    if (ksk) {   // EncryptionKeyChoice::Big: keyswitch first (gadget/engine/bootstrapping.rs:828-885)
        CK(ctx->ws.get(Workspace::B, (size_t)batch * small_len(p) * 8, &d_small));
        if (ks_dev_locked(ctx, ksk, d_in, d_small, batch)) return 1;
    }
    // bootstrap_common_factor :451-500 = blind rotation without sample extraction (entry a14)
    if (br_dev_locked(ctx, bsk, ksk ? d_small : d_in, d_glwe, d_lut, nullptr, batch, BR_MODE_GLWE)) return 1;
    int rc = glwe_mul_plain_extract_launch(d_glwe, d_pl, d_out, p.k, p.N, (int)n_polys, (int)batch, 64, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "glwe_mul_plain_extract launch failed: " + std::to_string(rc));
    CK(cudaMemcpyAsync(out, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int b200tfhe_pack_into_accumulator_batch_u64(b200tfhe_context *ctx, const b200tfhe_pksk *pksk, const uint64_t *lwes, uint32_t count,
                                             uint32_t p_mod, uint64_t *out_glwe, uint32_t batch) {
    REQUIRE(ctx && pksk, "pack_into_accumulator: null argument");
    REQUIRE(pksk->inner->ctx == ctx, "key handle belongs to another context");
    const Params &p = pksk->p;
    // the reference asserts an odd output modulus (bootstrapping.rs:719); every slot must be at least two coefficients wide
    REQUIRE(p_mod % 2 == 1 && p_mod >= 1 && (uint32_t)p.N / p_mod >= 2, "pack_into_accumulator: p must be odd with N / p >= 2");
    REQUIRE(count >= 1 && count <= p_mod, "pack_into_accumulator: between 1 and p ciphertexts per accumulator");
    if (batch == 0) return 0;
    REQUIRE(lwes && out_glwe, "pack_into_accumulator: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t len = big_len(p);
    const size_t in_b = (size_t)batch * count * len * 8, el_b = (size_t)batch * p_mod * len * 8;
    const size_t g_b = (size_t)batch * p_mod * glwe_len(p) * 8, out_b = (size_t)batch * glwe_len(p) * 8;
    void *d_in, *d_el, *d_g, *d_out;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::A, el_b, &d_el));
    CK(ctx->ws.get(Workspace::MID, g_b, &d_g));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, lwes, in_b, cudaMemcpyHostToDevice, ctx->stream));
    int rc = pack_elements_launch(d_in, d_el, (int)count, (int)p_mod, len, (int)batch, 64, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "pack_elements launch failed: " + std::to_string(rc));
    if (ks_dev_locked(ctx, pksk->inner, d_el, d_g, batch * p_mod)) return 1;
    rc = pack_rotate_sum_launch(d_g, d_out, (int)p_mod, p.k, p.N, (int)batch, 64, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "pack_rotate_sum launch failed: " + std::to_string(rc));
    CK(cudaMemcpyAsync(out_glwe, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// extract_bits (fft64/crypto/wop_pbs/mod.rs:66-225) over a batch: nbits rounds of {shift, keyswitch, store, bootstrap with a
// constant accumulator, subtract}, all device-resident; the ciphertexts of the batch advance through the rounds together.
int b200tfhe_extract_bits_batch_u64(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const uint64_t *in,
                                    uint32_t delta_log, uint32_t nbits, uint64_t *out, uint32_t batch) {
    REQUIRE(ctx, "extract_bits: null context");
    if (check_pair(ctx, ksk, bsk)) return 1;
    const Params &p = bsk->p;
    REQUIRE(p.bits == 64, "extract_bits: 64-bit torus only");
    // debug_assert of the reference (:86-94): the extracted window has to fit the scalar
    REQUIRE((uint64_t)delta_log + nbits <= 64, "extract_bits: delta_log + number_of_bits_to_extract exceeds the 64 bits of the ciphertext");
    // round bit_idx >= 1 uses alpha = 2^(delta_log + bit_idx - 1), round 0 with more than one bit 2^(delta_log - 1)   (:204)
    REQUIRE(nbits <= 1 || delta_log >= 1, "extract_bits: delta_log must be at least 1 to extract more than one bit");
    if (batch == 0 || nbits == 0) return 0;
    REQUIRE(in && out, "extract_bits: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t blen = big_len(p), slen = small_len(p), glen = glwe_len(p);
    const size_t in_b = (size_t)batch * blen * 8, out_b = (size_t)batch * nbits * slen * 8;
    void *d_buf, *d_shift, *d_ks, *d_pbs, *d_out, *d_lut;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_buf));
    CK(ctx->ws.get(Workspace::A, in_b, &d_shift));
    CK(ctx->ws.get(Workspace::B, (size_t)batch * slen * 8, &d_ks));
    CK(ctx->ws.get(Workspace::MID, in_b, &d_pbs));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(ctx->ws.get(Workspace::LUTS, (size_t)nbits * glen * 8, &d_lut));
    CK(cudaMemcpyAsync(d_buf, in, in_b, cudaMemcpyHostToDevice, ctx->stream));
    if (nbits > 1) {   // accumulator of round bit_idx: zero mask, body constant -2^(delta_log - 1 + bit_idx)   (:196-205)
        std::vector<uint64_t> luts((size_t)(nbits - 1) * glen, 0);
        for (uint32_t r = 0; r + 1 < nbits; r++)
            for (int c = 0; c < p.N; c++) luts[(size_t)r * glen + (size_t)p.k * p.N + c] = 0 - (1ULL << (delta_log - 1 + r));
        CK(cudaMemcpyAsync(d_lut, luts.data(), luts.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));   // luts is a local
    }
    for (uint32_t r = 0; r < nbits; r++) {
        const int sh = (int)(64 - delta_log - r - 1);
        const uint64_t alpha = r ? 1ULL << (delta_log + r - 2) : 0;   // added to the body of the PREVIOUS round's bootstrap (:219)
        int rc = eb_update_shift_launch(d_buf, r ? d_pbs : nullptr, d_shift, blen, (int)batch, sh, alpha, 64, ctx->stream);
        ctx->launches++;
        if (rc) return fail(ctx, "extract_bits shift launch failed: " + std::to_string(rc));
        if (ks_dev_locked(ctx, ksk, d_shift, d_ks, batch)) return 1;
        const bool last = r + 1 == nbits;
        rc = eb_store_launch(d_ks, d_out, slen, (int)batch, (int)nbits, (int)(nbits - 1 - r), last ? 0 : 1, 64, ctx->stream);
        ctx->launches++;
        if (rc) return fail(ctx, "extract_bits store launch failed: " + std::to_string(rc));
        if (last) break;
        if (br_dev_locked(ctx, bsk, d_ks, d_pbs, (const uint64_t *)d_lut + (size_t)r * glen, nullptr, batch, BR_MODE_LWE)) return 1;
    }
    CK(cudaMemcpyAsync(out, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---- circuit bootstrap + vertical packing, the second half of the WoP-PBS ---------------------------------------------
// LwePrivateFunctionalPackingKeyswitchKeyList of the circuit bootstrap (k+1 keys, algorithms/lwe_wopbs.rs:80-160) as k+1
// byte-plane matrices for the tensor-core keyswitch: [kN+1 input elements incl. the body][level * sub rows] x (k+1)N columns.
struct b200tfhe_pfpksk {
    b200tfhe_context *ctx;
    Params p;
    int base_log, level, sub, sub_bits;
    std::vector<KSTensorKey> keys;
    int device = 0;
};

int b200tfhe_cbs_pfpksk_upload_u64(b200tfhe_context *ctx, const b200tfhe_params *pp, uint32_t pfks_base_log, uint32_t pfks_level,
                                   const uint64_t *list, b200tfhe_pfpksk **out) {
    REQUIRE(ctx && pp && out && list, "cbs_pfpksk_upload: null argument");
    *out = nullptr;
    REQUIRE(pp->torus_bits == 64, "cbs_pfpksk_upload: 64-bit torus only");
    REQUIRE(pp->glwe_dimension > 0 && is_pow2(pp->polynomial_size), "invalid dimensions");
    REQUIRE(pfks_level >= 1 && pfks_base_log >= 1 && pfks_base_log <= 28 && pfks_level * pfks_base_log < 64, "invalid pfks decomposition");
    Params p = to_params(pp);
    const int sub = ((int)pfks_base_log + 5) / 6, sub_bits = ((int)pfks_base_log + sub - 1) / sub;   // sub-digits of <= 6 bits (+ carry) fit int8
    const int n_in = p.k * p.N + 1, n_out = (p.k + 1) * p.N - 1, rows_per = (int)pfks_level * sub;
    REQUIRE(ks_tc_supported(64, n_in, n_out, rows_per, sub_bits + 1), "cbs_pfpksk_upload: the tensor-core keyswitch does not cover this set");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t width = (size_t)n_out + 1, key_words = (size_t)n_in * pfks_level * width;
    void *d_raw = nullptr, *d_exp = nullptr;
    CK(cudaMalloc(&d_raw, key_words * 8));
    cudaError_t e = cudaMalloc(&d_exp, key_words * sub * 8);
    if (e != cudaSuccess) { cudaFree(d_raw); return fail_cuda(ctx, "cudaMalloc(pfpksk)", e); }
    b200tfhe_pfpksk *h = new b200tfhe_pfpksk{ctx, p, (int)pfks_base_log, (int)pfks_level, sub, sub_bits, {}};
    h->device = ctx->device;
    h->keys.resize(p.k + 1);
    int rc = 0;
    for (int r = 0; r <= p.k && !rc; r++) {
        void *planes = nullptr;
        if (cudaMemcpy(d_raw, list + (size_t)r * key_words, key_words * 8, cudaMemcpyHostToDevice) != cudaSuccess) { rc = -10; break; }
        if (cudaMalloc(&planes, ks_tc_key_bytes(64, n_in, n_out, rows_per)) != cudaSuccess) { rc = -11; break; }
        h->keys[r].d_planes = planes;
        // reference block order is level 1..L, consumed reversed (lwe_private_functional_packing_keyswitch.rs:81): rev = 1
        rc = ks_expand_rows_launch(d_raw, d_exp, (size_t)n_in, (int)pfks_level, sub, sub_bits, 1, width, 64, ctx->stream);
        if (!rc) rc = ks_tc_prepare_key(&h->keys[r], d_exp, planes, 64, n_in, n_out, rows_per, ctx->stream);
        if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = -12;
    }
    cudaFree(d_raw);
    cudaFree(d_exp);
    if (rc) {
        for (auto &k : h->keys) if (k.d_planes) cudaFree(k.d_planes);
        delete h;
        return fail(ctx, "cbs_pfpksk_upload failed: " + std::to_string(rc));
    }
    *out = h;
    return 0;
}
int b200tfhe_cbs_pfpksk_destroy(b200tfhe_pfpksk *k) {
    if (!k) return 0;
    cudaSetDevice(k->device);
    for (auto &t : k->keys) if (t.d_planes) cudaFree(t.d_planes);
    delete k;
    return 0;
}

// private_functional_keyswitch_lwe_ciphertext_into_glwe_ciphertext for key r over `entries` inputs (caller holds ctx->mu)
static int pfpks_dev_locked(b200tfhe_context *ctx, const b200tfhe_pfpksk *pf, int r, const void *d_in, void *d_out, size_t out_stride,
                            uint32_t entries) {
    const Params &p = pf->p;
    KSLaunch a{d_in, big_len(p), d_out, out_stride, nullptr, nullptr, p.k * p.N + 1, (p.k + 1) * p.N - 1, pf->level, pf->base_log, 64, (int)entries};
    a.sub = pf->sub;
    a.sub_bits = pf->sub_bits;
    a.no_body = true;
    void *digits = nullptr;
    CK(ctx->ws.get(Workspace::DIG, (size_t)entries * pf->keys[r].kp, &digits));
    int rc = ks_tc_launch(a, pf->keys[r], digits, ctx->stream);
    ctx->launches += 2;
    if (rc) return fail(ctx, "private functional packing keyswitch launch failed: " + std::to_string(rc));
    return 0;
}

int b200tfhe_pfpks_batch_u64(b200tfhe_context *ctx, const b200tfhe_pfpksk *pf, uint32_t key_index, const uint64_t *in, uint64_t *out_glwe,
                             uint32_t batch) {
    REQUIRE(ctx && pf, "pfpks: null argument");
    REQUIRE(pf->ctx == ctx, "key handle belongs to another context");
    REQUIRE(key_index < pf->keys.size(), "pfpks: key index out of range");
    if (batch == 0) return 0;
    REQUIRE(in && out_glwe, "pfpks: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const Params &p = pf->p;
    const size_t in_b = (size_t)batch * big_len(p) * 8, out_b = (size_t)batch * glwe_len(p) * 8;
    void *d_in, *d_out;
    CK(ctx->ws.get(Workspace::IN, in_b, &d_in));
    CK(ctx->ws.get(Workspace::OUT, out_b, &d_out));
    CK(cudaMemcpyAsync(d_in, in, in_b, cudaMemcpyHostToDevice, ctx->stream));
    if (pfpks_dev_locked(ctx, pf, (int)key_index, d_in, d_out, glwe_len(p), batch)) return 1;
    CK(cudaMemcpyAsync(out_glwe, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// circuit_bootstrap_boolean_vertical_packing (wop_pbs/mod.rs:647-853) over a batch.
int b200tfhe_cbs_vertical_packing_batch_u64(b200tfhe_context *ctx, const b200tfhe_bsk *bsk, const b200tfhe_pfpksk *pf, const uint64_t *bits_in,
                                            uint32_t nbits, const uint64_t *luts, uint32_t n_luts, uint32_t polys_per_lut,
                                            uint32_t cbs_base_log, uint32_t cbs_level, uint64_t *out, uint32_t batch) {
    REQUIRE(ctx && bsk && pf, "cbs_vertical_packing: null argument");
    REQUIRE(bsk->ctx == ctx && pf->ctx == ctx, "key handle belongs to another context");
    const Params &p = bsk->p;
    REQUIRE(p.bits == 64, "cbs_vertical_packing: 64-bit torus only");
    // the reference's debug_asserts (:277-309): fpksk input key = bootstrap output key, same GLWE shape
    REQUIRE(pf->p.k == p.k && pf->p.N == p.N, "cbs_vertical_packing: pfpksk and bootstrap key dimensions differ");
    REQUIRE(cbs_level >= 1 && cbs_base_log >= 1 && cbs_level * cbs_base_log < 64, "cbs_vertical_packing: invalid cbs decomposition");
    REQUIRE(nbits >= 1, "cbs_vertical_packing: empty lwe_list_in");   // :672-675
    int log2N = 0;
    while ((1 << log2N) < p.N) log2N++;
    // cmux_tree_memory_optimized asserts polynomial_count == 2^(number of its GGSWs) (:475); vertical_packing gives it the
    // log2(polynomial_count) most significant bits and the blind rotation the others (:812-825)
    REQUIRE(polys_per_lut >= 1 && (polys_per_lut & (polys_per_lut - 1)) == 0, "cbs_vertical_packing: polynomials per table must be a power of two");
    int tree = 0;
    while ((1u << tree) < polys_per_lut) tree++;
    REQUIRE(tree <= (int)nbits, "cbs_vertical_packing: more table polynomials than 2^nbits");
    const int n_br = (int)nbits - tree;
    REQUIRE(n_br <= log2N, "cbs_vertical_packing: the table is too small for the input bits (nbits - log2(polynomials per table) > log2 N)");
    Params pc = p;
    pc.pbs_level = (int)cbs_level;
    pc.pbs_base_log = (int)cbs_base_log;
    const int Rc = br_supported(pc);
    REQUIRE(Rc != 0 && Rc != 256, "cbs_vertical_packing: no blind-rotation instance for (N, k, cbs_level, cbs_base_log)");
    REQUIRE(tree == 0 || cmux_layer_supported(pc), "cbs_vertical_packing: no CMUX-tree instance for (N, k, cbs_level, cbs_base_log)");
    if (batch == 0 || n_luts == 0) return 0;
    REQUIRE(bits_in && luts && out, "cbs_vertical_packing: null buffer");
    // device scratch per input: nbits * cbs_level GGSW level matrices twice (standard + Fourier domain), 2 * 12 * 64 KB at the 2_2
    // WoP-PBS set; larger batches go through in slices so that the workspace stays around 2 GB
    constexpr uint32_t CBS_CHUNK = 1024;
    if (batch > CBS_CHUNK) {
        for (uint32_t b0 = 0; b0 < batch; b0 += CBS_CHUNK) {
            const uint32_t nb = batch - b0 < CBS_CHUNK ? batch - b0 : CBS_CHUNK;
            if (b200tfhe_cbs_vertical_packing_batch_u64(ctx, bsk, pf, bits_in + (size_t)b0 * nbits * small_len(p), nbits, luts, n_luts, polys_per_lut,
                                                        cbs_base_log, cbs_level, out + (size_t)b0 * n_luts * big_len(p), nb))
                return 1;
        }
        return 0;
    }
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t slen = small_len(p), blen = big_len(p), glen = glwe_len(p);
    const size_t entries = (size_t)batch * nbits * cbs_level;             // bootstraps = GGSW level matrices
    const size_t ggsw_words = (size_t)(p.k + 1) * glen;                   // one level matrix
    const size_t n_vp = (size_t)batch * n_luts;
    void *d_bits, *d_prep, *d_idx, *d_luts, *d_pbs, *d_ggsw, *d_fggsw, *d_out;
    CK(ctx->ws.get(Workspace::IN, (size_t)batch * nbits * slen * 8, &d_bits));
    CK(ctx->ws.get(Workspace::A, entries * slen * 8, &d_prep));
    CK(ctx->ws.get(Workspace::IDX, (entries + n_vp) * 4, &d_idx));
    const size_t n_direct = tree ? 0 : n_luts;                            // tables handed to the blind rotation as they are
    CK(ctx->ws.get(Workspace::LUTS, ((size_t)cbs_level + n_direct) * glen * 8 + (nbits + 1) * 8, &d_luts));
    CK(ctx->ws.get(Workspace::MID, entries * blen * 8, &d_pbs));
    CK(ctx->ws.get(Workspace::B, entries * ggsw_words * 8, &d_ggsw));
    CK(ctx->ws.get(Workspace::C, entries * ggsw_words * 8, &d_fggsw));
    CK(ctx->ws.get(Workspace::OUT, n_vp * blen * 8, &d_out));
    uint32_t *d_idx_vp = (uint32_t *)d_idx + entries;
    uint64_t *d_luts_vp = (uint64_t *)d_luts + (size_t)cbs_level * glen, *d_fake = d_luts_vp + n_direct * glen;
    // host tables: accumulators -2^(63 - b*level) (:423-427), the tables as trivial GLWEs, and the rotation amounts of
    // blind_rotate_assign (:872-890): step i divides by X^(2^i), i.e. a mask element whose modulus switch is 2N - 2^i, body 0
    std::vector<uint64_t> h(((size_t)cbs_level + n_direct) * glen + nbits + 1, 0);
    for (uint32_t l = 0; l < cbs_level; l++)
        for (int c = 0; c < p.N; c++) h[(size_t)l * glen + (size_t)p.k * p.N + c] = 0 - (1ULL << (63 - cbs_base_log * (l + 1)));
    for (size_t o = 0; o < n_direct; o++)
        memcpy(&h[((size_t)cbs_level + o) * glen + (size_t)p.k * p.N], luts + o * p.N, (size_t)p.N * 8);
    for (int i = 0; i < n_br; i++) h[((size_t)cbs_level + n_direct) * glen + i] = (uint64_t)(2 * p.N - (1 << i)) << (64 - log2N - 1);
    CK(cudaMemcpyAsync(d_luts, h.data(), h.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    std::vector<uint64_t> h_tab;                                          // CMUX tree leaves: every table polynomial as a trivial GLWE
    void *d_tab = nullptr;
    if (tree) {
        const size_t n_poly = (size_t)n_luts * polys_per_lut;
        h_tab.assign(n_poly * glen, 0);
        for (size_t q = 0; q < n_poly; q++) memcpy(&h_tab[q * glen + (size_t)p.k * p.N], luts + q * p.N, (size_t)p.N * 8);
        CK(ctx->ws.get(Workspace::D, h_tab.size() * 8, &d_tab));
        CK(cudaMemcpyAsync(d_tab, h_tab.data(), h_tab.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    }
    CK(cudaMemcpyAsync(d_bits, bits_in, (size_t)batch * nbits * slen * 8, cudaMemcpyHostToDevice, ctx->stream));
    int rc = cbs_prep_launch(d_bits, d_prep, (uint32_t *)d_idx, slen, (int)batch, (int)nbits, (int)cbs_level, 64, ctx->stream);
    // accumulator of rotation e: table e % n_luts as it is, or node e of the last tree layer
    if (!rc) rc = mod_index_launch(d_idx_vp, n_vp, tree ? (uint32_t)n_vp : n_luts, ctx->stream);
    ctx->launches += 2;
    if (rc) return fail(ctx, "cbs prep launch failed: " + std::to_string(rc));
    // homomorphic_shift_boolean for every (bit, level) in one bootstrap batch
    if (br_dev_locked(ctx, bsk, d_prep, d_pbs, d_luts, (const uint32_t *)d_idx, (uint32_t)entries, BR_MODE_LWE)) return 1;
    rc = cbs_post_launch(d_pbs, blen, entries, (int)cbs_level, (int)cbs_base_log, 64, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "cbs post launch failed: " + std::to_string(rc));
    // k+1 private functional packing keyswitches per bootstrap output: row r of the level matrix (:333-342)
    for (int r = 0; r <= p.k; r++)
        if (pfpks_dev_locked(ctx, pf, r, d_pbs, (uint64_t *)d_ggsw + (size_t)r * glen, ggsw_words, (uint32_t)entries)) return 1;
    // fill_with_forward_fourier (:732): every polynomial of every GGSW, device layout of the blind-rotation kernels
    rc = bsk_convert_polys_launch(Rc, 64, entries * (p.k + 1) * (p.k + 1), d_ggsw, d_fggsw, tw_fwd(ctx, Rc), ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "ggsw conversion launch failed: " + std::to_string(rc));
    const size_t ggsw_cplx = (size_t)cbs_level * (p.k + 1) * (p.k + 1) * (p.N / 2), key_stride = (size_t)nbits * ggsw_cplx;
    // cmux_tree_memory_optimized (:468-592), a layer per launch: layer j pairs consecutive nodes with the j-th GGSW from the end of
    // the tree's sub-list = entry n_br + j of the reversed list built in step 1
    const void *d_acc = d_luts_vp;
    if (tree) {
        // the conversion above was the last reader of the standard-domain GGSWs and of the bootstrap inputs: their slots ping-pong
        void *d_layer[2];
        const size_t layer_bytes = n_vp * (polys_per_lut / 2) * glen * 8;
        CK(ctx->ws.get(Workspace::A, layer_bytes, &d_layer[0]));
        CK(ctx->ws.get(Workspace::B, layer_bytes, &d_layer[1]));
        const void *src = d_tab;
        for (int j = 0; j < tree; j++) {
            const int pairs = (int)(polys_per_lut >> (j + 1));
            rc = cmux_layer_launch(pc, src, j ? (size_t)n_luts * pairs * 2 : 0, d_layer[j & 1], (const cplx *)d_fggsw + (size_t)(n_br + j) * ggsw_cplx,
                                   key_stride, (int)n_luts, pairs, n_vp * pairs, tw_fwd(ctx, Rc), tw_inv(ctx, Rc), ctx->stream);
            ctx->launches++;
            if (rc) return fail(ctx, "cmux tree launch failed: " + std::to_string(rc));
            src = d_layer[j & 1];
        }
        d_acc = src;
    }
    // vertical_packing (:812-853): blind rotation over the ciphertext's own remaining GGSWs + sample extraction
    BRLaunch a{d_fake, 0, d_acc, d_idx_vp, d_fggsw, d_out, blen, n_br, (int)n_vp, BR_MODE_LWE, tw_fwd(ctx, Rc), tw_inv(ctx, Rc)};
    a.bsk_ct_stride = key_stride;
    a.bsk_ct_div = (int)n_luts;
    rc = br_launch(pc, a, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "vertical packing launch failed: " + std::to_string(rc));
    CK(cudaMemcpyAsync(out, d_out, n_vp * blen * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---- device-resident waves over a ciphertext pool -------------------------------------------------------------------
static_assert(sizeof(b200tfhe_wave_op) == sizeof(WaveOp) && sizeof(WaveOp) == 72, "wave op layout");
struct b200tfhe_pool {
    b200tfhe_context *ctx;
    Params p;
    bool big;
    uint32_t capacity;
    size_t len;                       // scalars per slot
    void *d_slots = nullptr, *d_luts = nullptr, *d_ops = nullptr, *d_staging = nullptr, *d_mid = nullptr, *d_out = nullptr;
    uint32_t *d_idx = nullptr;
    uint32_t n_luts = 0, ops_cap = 0;
    int device = 0;
};

int b200tfhe_pool_create(b200tfhe_context *ctx, const b200tfhe_params *pp, int big, uint32_t capacity, b200tfhe_pool **out) {
    REQUIRE(ctx && pp && out, "pool_create: null argument");
    *out = nullptr;
    REQUIRE(capacity > 0, "pool_create: empty pool");
    Params p = to_params(pp);
    if (check_pbs_params(ctx, p) || check_ks_params(ctx, p)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    b200tfhe_pool *pl = new b200tfhe_pool{ctx, p, big != 0, capacity, big ? big_len(p) : small_len(p)};
    pl->device = ctx->device;
    cudaError_t e = cudaMalloc(&pl->d_slots, (size_t)capacity * pl->len * sz(p));
    if (e != cudaSuccess) { delete pl; return fail_cuda(ctx, "cudaMalloc(pool)", e); }
    cudaMemsetAsync(pl->d_slots, 0, (size_t)capacity * pl->len * sz(p), ctx->stream);
    *out = pl;
    return 0;
}
int b200tfhe_pool_destroy(b200tfhe_pool *pl) {
    if (!pl) return 0;
    cudaSetDevice(pl->device);
    cudaDeviceSynchronize();     // not the context's stream: the context may already be destroyed
    for (void *q : {pl->d_slots, pl->d_luts, pl->d_ops, pl->d_staging, pl->d_mid, pl->d_out, (void *)pl->d_idx})
        if (q) cudaFree(q);
    delete pl;
    return 0;
}
int b200tfhe_pool_write(b200tfhe_context *ctx, b200tfhe_pool *pl, uint32_t first, uint32_t count, const void *host) {
    REQUIRE(ctx && pl && pl->ctx == ctx, "pool_write: bad handle");
    REQUIRE((uint64_t)first + count <= pl->capacity, "pool_write: slot range out of bounds");
    if (count == 0) return 0;
    REQUIRE(host, "pool_write: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t bytes = pl->len * sz(pl->p);
    CK(cudaMemcpyAsync((char *)pl->d_slots + first * bytes, host, count * bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int b200tfhe_pool_read(b200tfhe_context *ctx, const b200tfhe_pool *pl, uint32_t first, uint32_t count, void *host) {
    REQUIRE(ctx && pl && pl->ctx == ctx, "pool_read: bad handle");
    REQUIRE((uint64_t)first + count <= pl->capacity, "pool_read: slot range out of bounds");
    if (count == 0) return 0;
    REQUIRE(host, "pool_read: null buffer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    const size_t bytes = pl->len * sz(pl->p);
    CK(cudaMemcpyAsync(host, (const char *)pl->d_slots + first * bytes, count * bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int b200tfhe_pool_device_buffer(const b200tfhe_pool *pl, void **d_ptr, size_t *bytes) {
    if (!pl || !d_ptr || !bytes) return 1;
    *d_ptr = pl->d_slots;
    *bytes = (size_t)pl->capacity * pl->len * sz(pl->p);
    return 0;
}
int b200tfhe_pool_set_luts(b200tfhe_context *ctx, b200tfhe_pool *pl, const void *host_luts, uint32_t n_luts) {
    REQUIRE(ctx && pl && pl->ctx == ctx, "pool_set_luts: bad handle");
    REQUIRE(host_luts && n_luts > 0, "pool_set_luts: null / empty accumulator list");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (pl->d_luts) { cudaFree(pl->d_luts); pl->d_luts = nullptr; }
    const size_t bytes = (size_t)n_luts * glwe_len(pl->p) * sz(pl->p);
    CK(cudaMalloc(&pl->d_luts, bytes));
    CK(cudaMemcpy(pl->d_luts, host_luts, bytes, cudaMemcpyHostToDevice));
    pl->n_luts = n_luts;
    return 0;
}
// argument checks of one wave (no device work)
static int wave_check(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, const b200tfhe_pool *pl,
                      const b200tfhe_wave_op *ops, uint32_t n_ops, uint32_t n_pbs) {
    REQUIRE(ops && n_pbs <= n_ops, "wave_run: bad operation list");
    if (n_pbs) {
        if (check_pair(ctx, ksk, bsk)) return 1;
        REQUIRE(pl->n_luts > 0, "wave_run: the pool has no accumulators (b200tfhe_pool_set_luts)");
        const Params &p = pl->p, &kp = bsk->p;
        REQUIRE(kp.n == p.n && kp.k == p.k && kp.N == p.N && kp.bits == p.bits, "wave_run: keys and pool have different parameters");
    }
    for (uint32_t i = 0; i < n_ops; i++) {
        const b200tfhe_wave_op &o = ops[i];
        REQUIRE(o.dst < pl->capacity, "wave_run: destination slot out of range");
        for (int k = 0; k < 4; k++) REQUIRE(o.scale[k] == 0 || o.src[k] < pl->capacity, "wave_run: source slot out of range");
        REQUIRE(o.scale4 == 0 || o.src4 < pl->capacity, "wave_run: source slot out of range");
        if (i < n_pbs) REQUIRE(o.lut >= 0 && (uint32_t)o.lut < pl->n_luts, "wave_run: accumulator index out of range");
        else REQUIRE(o.lut < 0, "wave_run: bootstrapped operations must come first");
    }
    return 0;
}
// one wave on the context's stream (caller holds ctx->mu, device is current, arguments are checked)
static int wave_run_locked(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, b200tfhe_pool *pl,
                           const b200tfhe_wave_op *ops, uint32_t n_ops, uint32_t n_pbs) {
    const Params &p = pl->p;
    const size_t s = sz(p), len = pl->len, mid_len = pl->big ? small_len(p) : big_len(p);
    if (pl->ops_cap < n_ops) {
        CK(cudaStreamSynchronize(ctx->stream));
        for (void *q : {pl->d_ops, pl->d_staging, pl->d_mid, pl->d_out, (void *)pl->d_idx})
            if (q) cudaFree(q);
        pl->d_ops = pl->d_staging = pl->d_mid = pl->d_out = nullptr; pl->d_idx = nullptr; pl->ops_cap = 0;
        uint32_t cap = n_ops < 256 ? 256 : n_ops;
        CK(cudaMalloc(&pl->d_ops, (size_t)cap * sizeof(WaveOp)));
        CK(cudaMalloc(&pl->d_staging, (size_t)cap * len * s));
        CK(cudaMalloc(&pl->d_out, (size_t)cap * len * s));
        CK(cudaMalloc(&pl->d_mid, (size_t)cap * mid_len * s));
        CK(cudaMalloc((void **)&pl->d_idx, (size_t)cap * 4));
        pl->ops_cap = cap;
    }
    // pageable source: cudaMemcpyAsync returns once the operations are staged, so the caller's array can be reused; the copy is
    // stream-ordered after the previous wave's kernels, which still read d_ops
    CK(cudaMemcpyAsync(pl->d_ops, ops, (size_t)n_ops * sizeof(WaveOp), cudaMemcpyHostToDevice, ctx->stream));
    int rc = wave_gather_launch(pl->d_slots, len, (const WaveOp *)pl->d_ops, (int)n_ops, pl->d_staging, pl->d_idx, (int)n_pbs, p.bits, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "wave gather launch failed: " + std::to_string(rc));
    if (n_pbs) {
        if (pl->big) {
            if (ks_dev_locked(ctx, ksk, pl->d_staging, pl->d_mid, n_pbs)) return 1;
            if (br_dev_locked(ctx, bsk, pl->d_mid, pl->d_out, pl->d_luts, pl->d_idx, n_pbs, BR_MODE_LWE)) return 1;
        } else {
            if (br_dev_locked(ctx, bsk, pl->d_staging, pl->d_mid, pl->d_luts, pl->d_idx, n_pbs, BR_MODE_LWE)) return 1;
            if (ks_dev_locked(ctx, ksk, pl->d_mid, pl->d_out, n_pbs)) return 1;
        }
    }
    rc = wave_scatter_launch(pl->d_slots, len, (const WaveOp *)pl->d_ops, (int)n_ops, (int)n_pbs, pl->d_out, pl->d_staging, p.bits, ctx->stream);
    ctx->launches++;
    if (rc) return fail(ctx, "wave scatter launch failed: " + std::to_string(rc));
    return 0;
}
int b200tfhe_wave_run(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, b200tfhe_pool *pl,
                      const b200tfhe_wave_op *ops, uint32_t n_ops, uint32_t n_pbs) {
    REQUIRE(ctx && pl && pl->ctx == ctx, "wave_run: bad handle");
    if (n_ops == 0) return 0;
    if (wave_check(ctx, ksk, bsk, pl, ops, n_ops, n_pbs)) return 1;
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    return wave_run_locked(ctx, ksk, bsk, pl, ops, n_ops, n_pbs);
}
// A whole pre-scheduled circuit in ONE call: the waves of `ops` (concatenated; wave w has n_ops[w] operations, the first
// n_pbs[w] of them bootstrapped) run back to back on the context's stream -- the host-side scheduler (Python OpGraph, or
// a Rust caller's rayon layers) is off the critical path: a circuit is compiled once and replayed.
int b200tfhe_wave_sequence_run(b200tfhe_context *ctx, const b200tfhe_ksk *ksk, const b200tfhe_bsk *bsk, b200tfhe_pool *pl,
                               const b200tfhe_wave_op *ops, const uint32_t *n_ops, const uint32_t *n_pbs, uint32_t n_waves) {
    REQUIRE(ctx && pl && pl->ctx == ctx, "wave_sequence_run: bad handle");
    if (n_waves == 0) return 0;
    REQUIRE(ops && n_ops && n_pbs, "wave_sequence_run: null argument");
    size_t off = 0;
    for (uint32_t w = 0; w < n_waves; w++) {
        if (n_ops[w] && wave_check(ctx, ksk, bsk, pl, ops + off, n_ops[w], n_pbs[w])) return 1;
        off += n_ops[w];
    }
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    off = 0;
    for (uint32_t w = 0; w < n_waves; w++) {
        if (n_ops[w] && wave_run_locked(ctx, ksk, bsk, pl, ops + off, n_ops[w], n_pbs[w])) return 1;
        off += n_ops[w];
    }
    return 0;
}

// Measures the sustained FP64 FMA rate of the device (TFLOP/s, 2 flop per FMA) with a register-only kernel:
// the denominator of the blind rotation's FP64 roofline.  duration_ms: how long to run (>= 20).
int b200tfhe_probe_fp64_tflops(b200tfhe_context *ctx, double duration_ms, double *tflops) {
    REQUIRE(ctx && tflops, "probe: null argument");
    std::lock_guard<std::mutex> g(ctx->mu);
    CK(cudaSetDevice(ctx->device));
    void *scratch;
    CK(ctx->ws.get(Workspace::A, 64, &scratch));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    const int grid = sms * 8, iters = 4096;
    const double flop_per_launch = 2.0 * 64.0 * iters * 256.0 * grid;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int w = 0; w < 3; w++) dfma_probe_launch((double *)scratch, grid, iters, ctx->stream);
    double best = 0.0, spent = 0.0;
    while (spent < duration_ms) {
        CK(cudaEventRecord(e0, ctx->stream));
        for (int r = 0; r < 4; r++) dfma_probe_launch((double *)scratch, grid, iters, ctx->stream);
        CK(cudaEventRecord(e1, ctx->stream));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ctx->launches += 4;
        spent += ms;
        double tf = 4.0 * flop_per_launch / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return 0;
}

}  // extern "C"

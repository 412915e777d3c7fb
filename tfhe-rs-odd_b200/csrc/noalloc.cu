// noalloc.cu -- the allocation-free keyswitch variant (b200tfhe_keyswitch_batch_noalloc_dev, include/b200tfhe.h).
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The tensor-core keyswitch (keyswitch_tc.cu) needs a
// scratch matrix of int8 digits, batch x padded(kN * l_ks) bytes, which the ordinary entry points take from the context's
// growable workspace (cudaMalloc on first use / on growth).  Here that scratch is a file-scope device array of fixed size:
// every GPU that loads the library holds NOALLOC_KS_SCRATCH_BYTES of it until the process exits.
//
// In host code `g_noalloc_ks_digits` is only a host-side stand-in for the symbol, NOT a device address: the device address is
// looked up once per device with cudaGetSymbolAddress (with that device current) and it is that pointer the kernels get.
// All calls (from every context on the device) share the one array, and it keeps the previous call's contents: nothing in it
// has to start at zero -- ks_digits_kernel rewrites every byte the GEMM reads, and the tensor map's extent excludes the row
// padding -- but two calls that overlap on the GPU (different contexts / streams on one device) race on it.
#include "kernels.h"

namespace b200tfhe {

static __device__ __align__(256) unsigned char g_noalloc_ks_digits[NOALLOC_KS_SCRATCH_BYTES];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
void *g_scratch_addr[MAX_DEVICES] = {};   // device address of g_noalloc_ks_digits on each device, filled on first use
}  // namespace

// Caller has made `device` current.  Returns 0 and launches ks_digits_kernel + ks_tc_kernel on `st`, or a negative code:
// -10 scratch too small for this batch, -11 device index out of range, -12 symbol lookup failed (*cuda_err says why).
int ks_tc_noalloc_launch(const KSLaunch &a, const KSTensorKey &tk, int device, cudaStream_t st, cudaError_t *cuda_err) {
    *cuda_err = cudaSuccess;
    if ((size_t)a.batch * tk.kp > NOALLOC_KS_SCRATCH_BYTES) return -10;
    if (device < 0 || device >= MAX_DEVICES) return -11;
    void *ptr = nullptr;
    {
        std::lock_guard<std::mutex> g(g_mu);
        if (!g_scratch_addr[device]) {
            cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_ks_digits);
            if (e != cudaSuccess || !ptr) {
                *cuda_err = e;
                return -12;
            }
            g_scratch_addr[device] = ptr;
        }
        ptr = g_scratch_addr[device];
    }
    return ks_tc_launch(a, tk, ptr, st);
}

}  // namespace b200tfhe

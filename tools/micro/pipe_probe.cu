// pipe_probe.cu -- the DMMA-vs-DFMA experiment BASELINE.json's north_star asks for, plus two conversions.
// Measures on one B200 (148 SMs, 16 warps per SM, timed with CUDA events):
//   dfma        8 independent DFMA chains per thread                            -> TFLOP/s of the vector FP64 pipe
//   dmma        4 independent mma.sync.m8n8k4.f64 accumulators per warp         -> TFLOP/s of the FP64 tensor path
//   dfma+dmma   both interleaved in the same warp                               -> do the two paths add up (separate pipes)?
//   split       half of the warps DFMA only, half DMMA only                     -> same question, warp-level overlap
//   f2i         cvt.rni.s64.f64 (the rounding of torus_from_double)             -> conversions per clock per SM
//   magic       the conversion-free alternative: 2 DADD + integer ops            -> same unit
// Build/run:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu && ./pipe_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// MODE 0: dfma, 1: dmma, 2: both in one warp, 3: split by warp parity
template <int MODE> __global__ void __launch_bounds__(512) fp_probe(double *out, double x, double y, int iters) {
    double f[8], c[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { f[j] = threadIdx.x + j; c[j] = j; }
    const bool do_f = MODE == 0 || MODE == 2 || (MODE == 3 && ((threadIdx.x >> 5) & 1) == 0);
    const bool do_m = MODE == 1 || MODE == 2 || (MODE == 3 && ((threadIdx.x >> 5) & 1) == 1);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 4; rep++) {
            if (do_f) {
#pragma unroll
                for (int j = 0; j < 8; j++) f[j] = fma(f[j], x, y);
            }
            if (do_m) {
#pragma unroll
                for (int j = 0; j < 4; j++) dmma(c[2 * j], c[2 * j + 1], x, y);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += f[j] + c[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// MODE 0: F2I.S64.F64 of (t - rint(t)) * 2^64 as torus_from_double does it; MODE 1: the same value through two magic adds
template <int MODE> __global__ void __launch_bounds__(512) cvt_probe(unsigned long long *out, double x, int iters) {
    double t[8];
    unsigned long long acc = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) t[j] = (threadIdx.x + 1) * 0.37 + j * 1.13;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (MODE == 0) {
                const double MAGIC = 6755399441055744.0;
                double r = (t[j] + MAGIC) - MAGIC, fr = (t[j] - r) * 18446744073709551616.0;
                acc += (unsigned long long)__double2ll_rn(fr);
            } else {
                // s = t * 2^64 (the scale is free: it folds into the inverse twiddle table); hi part on a 2^44 grid, low part exact
                const double s = t[j] * 18446744073709551616.0;
                const double M1 = 6755399441055744.0 * 17592186044416.0;   // 1.5 * 2^52 * 2^44
                const double a1 = s + M1, r1 = s - (a1 - M1);
                const double a2 = r1 + 6755399441055744.0;
                const unsigned long long hi = (unsigned long long)(unsigned)__double2loint(a1) << 44;
                const long long lo = ((long long)__double_as_longlong(a2) << 13) >> 13;
                acc += hi + (unsigned long long)lo;
            }
            t[j] = t[j] * x + 0.001;   // keeps the inputs moving (one DFMA per conversion in both modes)
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <class F> float time_ms(F launch) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    double *d; unsigned long long *u;
    cudaMalloc(&d, 148 * 4 * 512 * 8); cudaMalloc(&u, 148 * 4 * 512 * 8);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int iters = 20000, grid = 148, threads = 512;
    const double warps = (double)grid * threads / 32.0;
    const double dfma_flop = warps * 32.0 * iters * 4.0 * 8.0 * 2.0;            // per launch, MODE 0
    const double dmma_flop = warps * iters * 4.0 * 4.0 * (8.0 * 8.0 * 4.0 * 2.0);   // per launch, MODE 1
    float t0 = time_ms([&] { fp_probe<0><<<grid, threads>>>(d, 1.0000001, 1e-9, iters); });
    float t1 = time_ms([&] { fp_probe<1><<<grid, threads>>>(d, 1.0000001, 1e-9, iters); });
    float t2 = time_ms([&] { fp_probe<2><<<grid, threads>>>(d, 1.0000001, 1e-9, iters); });
    float t3 = time_ms([&] { fp_probe<3><<<grid, threads>>>(d, 1.0000001, 1e-9, iters); });
    printf("{\"sm_clock_khz_max\": %d,\n", clk_khz);
    printf(" \"dfma_only\": {\"ms\": %.3f, \"tflops\": %.2f},\n", t0, dfma_flop / t0 * 1e-9);
    printf(" \"dmma_only\": {\"ms\": %.3f, \"tflops\": %.2f},\n", t1, dmma_flop / t1 * 1e-9);
    printf(" \"dfma_plus_dmma_same_warp\": {\"ms\": %.3f, \"tflops_total\": %.2f, \"ms_if_serial\": %.3f},\n", t2, (dfma_flop + dmma_flop) / t2 * 1e-9, t0 + t1);
    printf(" \"dfma_dmma_split_warps\": {\"ms\": %.3f, \"tflops_total\": %.2f, \"ms_if_serial\": %.3f},\n", t3, (dfma_flop + dmma_flop) / 2 / t3 * 1e-9, (t0 + t1) / 2);
    const double ncvt = warps * 32.0 * iters * 8.0;
    float c0 = time_ms([&] { cvt_probe<0><<<grid, threads>>>(u, 1.0000001, iters); });
    float c1 = time_ms([&] { cvt_probe<1><<<grid, threads>>>(u, 1.0000001, iters); });
    printf(" \"round_f2i\": {\"ms\": %.3f, \"gconv_per_s\": %.1f},\n", c0, ncvt / c0 * 1e-6);
    printf(" \"round_magic\": {\"ms\": %.3f, \"gconv_per_s\": %.1f}}\n", c1, ncvt / c1 * 1e-6);
    return 0;
}

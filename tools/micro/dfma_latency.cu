// dfma_latency.cu -- how many independent DFMA chains does one warp need to keep the B200 FP64 pipe busy?
// Prints cycles per DFMA (per warp) for ILP = 1..16 with 1, 2 and 4 warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dfma_latency dfma_latency.cu && ./dfma_latency
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void probe(double *out, unsigned long long *cycles, double x, double y, int iters) {
    double a[ILP];
#pragma unroll
    for (int j = 0; j < ILP; j++) a[j] = threadIdx.x + j;
    unsigned long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++)
#pragma unroll
            for (int j = 0; j < ILP; j++) a[j] = fma(a[j], x, y);
    }
    unsigned long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int j = 0; j < ILP; j++) s += a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int ILP> void run(double *d_out, unsigned long long *d_cyc, int threads) {
    const int iters = 2000;
    probe<ILP><<<148, threads>>>(d_out, d_cyc, 1.0000001, 1e-9, iters);
    cudaDeviceSynchronize();
    probe<ILP><<<148, threads>>>(d_out, d_cyc, 1.0000001, 1e-9, iters);
    unsigned long long c;
    cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("  ILP %2d: %.2f cycles per DFMA per warp\n", ILP, (double)c / (iters * 8.0 * ILP));
}

int main() {
    double *d_out; unsigned long long *d_cyc;
    cudaMalloc(&d_out, 148 * 1024 * 8); cudaMalloc(&d_cyc, 8);
    for (int wps : {1, 2, 4}) {
        printf("%d warp(s) per sub-partition:\n", wps);
        run<1>(d_out, d_cyc, wps * 128); run<2>(d_out, d_cyc, wps * 128); run<3>(d_out, d_cyc, wps * 128);
        run<4>(d_out, d_cyc, wps * 128); run<6>(d_out, d_cyc, wps * 128); run<8>(d_out, d_cyc, wps * 128); run<16>(d_out, d_cyc, wps * 128);
    }
    return 0;
}

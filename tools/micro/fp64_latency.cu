// Dependent-chain latency of the FP64 instructions the recursion's critical path is made of (one warp, one SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu && ./fp64_latency
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_lat(double* out, long long* cyc, int n, double a0, double b0)
{
    double a = a0 + threadIdx.x, b = b0;
    long long t0 = clock64();
    for (int i = 0; i < n; i++) { a = __dadd_rn(a, b); a = __dadd_rn(a, b); a = __dadd_rn(a, b); a = __dadd_rn(a, b); }
    long long t1 = clock64();
    double c = a0 + threadIdx.x, d = b0 * 0.5;
    for (int i = 0; i < n; i++) {       // compare + select chains (max)
        c = (c > d) ? c : d + 1.0; c = (c > d) ? c : d + 1.0; c = (c > d) ? c : d + 1.0; c = (c > d) ? c : d + 1.0;
    }
    long long t2 = clock64();
    double e = a0;
    for (int i = 0; i < n; i++) { e = __fma_rn(e, b, b); e = __fma_rn(e, b, b); e = __fma_rn(e, b, b); e = __fma_rn(e, b, b); }
    long long t3 = clock64();
    int k = threadIdx.x;
    for (int i = 0; i < n; i++) { k = (k > i) ? k - 1 : k + 3; k = (k > i) ? k - 1 : k + 3; k = (k > i) ? k - 1 : k + 3; k = (k > i) ? k - 1 : k + 3; }
    long long t4 = clock64();
    out[threadIdx.x] = a + c + e + k;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; }
}
int main()
{
    double* d_out; long long* d_cyc; long long h[4];
    cudaMalloc(&d_out, 32 * 8); cudaMalloc(&d_cyc, 32);
    const int n = 4096;
    for (int rep = 0; rep < 2; rep++) k_lat<<<1, 32>>>(d_out, d_cyc, n, 1.0, 1e-9);
    cudaMemcpy(h, d_cyc, 32, cudaMemcpyDeviceToHost);
    printf("DADD dependent: %.1f cycles/op\nDSETP+select(+DADD) dependent: %.1f cycles/iter\nDFMA dependent: %.1f cycles/op\nint cmp+sel+add dependent: %.1f cycles/iter\n",
           h[0] / (4.0 * n), h[1] / (4.0 * n), h[2] / (4.0 * n), h[3] / (4.0 * n));
    return 0;
}

// Cycles per step of the register-resident recursion (one warp), with the emissions read from shared memory
// and (variant 2) the byte store of the decoded state.   nvcc -I../../signals_b200/csrc ...
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "kernels.h"
template <int NL, int VARIANT>
__global__ void k_step(double* out, long long* cyc, uint8_t* path, int nsteps, double tr_self, double tr_off)
{
    __shared__ __align__(16) double sm_all[4][32 * 4 * 4 * (NL / 4)];
    double* sm = sm_all[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;            // 64 bins x 4 states per lane (NL = 4) -- reused cyclically
    for (int i = lane; i < 32 * 4 * 4 * (NL / 4); i += 32) sm[i] = -1.0 - (double)((i * 2654435761u) >> 20) * 1e-3;
    __syncwarp();
    double T1[NL], m1 = 0.0;
    for (int i = 0; i < NL; i++) T1[i] = -(double)i;
    const double* mine = sm + lane * 4;
    uint8_t* outp = path + ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * nsteps;
    unsigned acc = 0;
    long long t0 = clock64();
    for (int t = 0; t < nsteps; t += 4) {
#pragma unroll
        for (int b = 0; b < 4; b++) {
            double e[NL];
#pragma unroll
            for (int p = 0; p < NL / 4; p++) {
                const double* q;
                if (VARIANT & 4) {
                    constexpr int RUN = NL;                 // 4 * planes sectors per lane per chunk
                    int pos = b * (NL / 4) + p + (lane % RUN);
                    pos = pos >= RUN ? pos - RUN : pos;
                    q = sm + (lane * RUN + pos) * 4;
                } else q = mine + (((t + b) & 3) * 32 * 4) + p * 0;
                const double2 lo = *(const double2*)q, hi = *(const double2*)(q + 2);
                e[4 * p] = lo.x; e[4 * p + 1] = lo.y; e[4 * p + 2] = hi.x; e[4 * p + 3] = hi.y;
            }
            int back;
            if (VARIANT & 1) {
                bool tie;
                back = sgb_seq_step_fast<NL>(T1, e, m1, tr_self, tr_off, tie);
                acc |= tie ? 1u << b : 0u;
            } else {
                back = sgb_seq_step_r<NL>(T1, e, m1, tr_self, tr_off);
            }
            if (VARIANT & 2) outp[t + b] = (uint8_t)back; else acc += back;
        }
    }
    long long t1 = clock64();
    out[threadIdx.x] = m1 + acc;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int NL, int V>
static void run(const char* what, double* d_out, long long* d_cyc, uint8_t* d_path, int n, int warps = 1, int blocks = 1)
{
    long long h = 0;
    for (int r = 0; r < 2; r++) k_step<NL, V><<<blocks, 32 * warps>>>(d_out, d_cyc, d_path, n, -0.0513, -5.39);
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-40s warps %d blocks %3d NL=%2d: %.1f cycles/step  (%s)\n", what, warps, blocks, NL, (double)h / n, cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    double* d_out; long long* d_cyc; uint8_t* d_path;
    const int n = 4096;
    cudaMalloc(&d_out, 32 * 8); cudaMalloc(&d_cyc, 32); cudaMalloc(&d_path, (size_t)32 * n * 8 * 296);
    run<4, 0>("full step (tie branch), no store", d_out, d_cyc, d_path, n);
    run<4, 1>("branch-free step, no store", d_out, d_cyc, d_path, n);
    run<4, 2>("full step, byte store", d_out, d_cyc, d_path, n);
    run<4, 3>("branch-free step, byte store", d_out, d_cyc, d_path, n);
    run<8, 1>("branch-free step, no store", d_out, d_cyc, d_path, n);
    run<12, 1>("branch-free step, no store", d_out, d_cyc, d_path, n);
    run<12, 3>("branch-free step, byte store", d_out, d_cyc, d_path, n);
    run<12, 1>("branch-free, no store", d_out, d_cyc, d_path, n, 4, 1);
    run<12, 1>("branch-free, no store", d_out, d_cyc, d_path, n, 4, 148);
    run<12, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 1, 1);
    run<12, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 4, 148);
    run<4, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 1, 1);
    run<4, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 4, 148);
    run<4, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 4, 296);
    run<4, 1>("branch-free, no store", d_out, d_cyc, d_path, n, 4, 148);
    run<8, 5>("branch-free, no store, rotated layout", d_out, d_cyc, d_path, n, 4, 148);
    return 0;
}

// The fast chunk body of k_recur (4 bins) in isolation, features switched on one by one (FEAT bits):
//   1 ties collected + warp vote + redo branch   2 staged byte + lastback   4 T1s copy kept live   8 mbarrier-like fence (syncwarp) per chunk
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "kernels.h"
#define FULL 0xffffffffu
template <int NPL, int FEAT>
__global__ void __launch_bounds__(128, 2) k_chunk(double* out, long long* cyc, int nchunks, double tr_self, double tr_off, int len)
{
    constexpr int NL = 4 * NPL, RUN = 4 * NPL;
    extern __shared__ __align__(16) unsigned char smraw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned char* ring = smraw + wid * (NPL * 4096 + 1088);
    unsigned char* stage = ring + NPL * 4096;
    for (int i = lane; i < NPL * 512; i += 32) ((double*)ring)[i] = -1.0 - (double)((i * 2654435761u) >> 20) * 1e-3;
    __syncwarp();
    if (FEAT & 16) {          // states >= NL - 2 masked (-Inf) as in real data: position of state st of (lane l, bin b, plane p)
        for (int l = 0; l < 32; l++)
            for (int b = 0; b < 4; b++) {
                int pos = b * NPL + (NPL - 1) + (l % RUN);
                pos = pos >= RUN ? pos - RUN : pos;
                double* q = (double*)(ring + l * (RUN * 32) + pos * 32);
                if (lane == 0) { q[2] = -sgb_inf(); q[3] = -sgb_inf(); }
            }
    }
    __syncwarp();
    double T1[NL], m1 = 0.0;
    for (int i = 0; i < NL; i++) T1[i] = -(double)i;
    const int lrm = lane % RUN;
    const unsigned char* mine = ring + lane * (RUN * 32);
    int lastback = 0;
    unsigned acc = 0;
    long long t0 = clock64();
    for (int c = 1; c <= nchunks; c++) {
        auto load_e = [&](int b, double (&e)[NL]) {
#pragma unroll
            for (int p = 0; p < NPL; p++) {
                int pos = b * NPL + p + lrm;
                pos = pos >= RUN ? pos - RUN : pos;
                const double2 lo = *(const double2*)(mine + pos * 32), hi = *(const double2*)(mine + pos * 32 + 16);
                e[4 * p] = lo.x; e[4 * p + 1] = lo.y; e[4 * p + 2] = hi.x; e[4 * p + 3] = hi.y;
            }
        };
        double T1s[NL];
        if (FEAT & 4) {
#pragma unroll
            for (int i = 0; i < NL; i++) T1s[i] = T1[i];
        }
        const double m1s = m1;
        int back[4];
        unsigned ties = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            double e[NL];
            load_e(b, e);
            bool tie;
            back[b] = sgb_seq_step_fast<NL>(T1, e, m1, tr_self, tr_off, tie);
            if (FEAT & 1) ties |= (tie && 4 * c + b < len) ? 1u << b : 0u;
        }
        if (FEAT & 1) {
            if (__any_sync(FULL, ties != 0)) {
                if (ties) {
                    if (FEAT & 4) {
                        double m1r = m1s;
#pragma unroll 1
                        for (int b = 0; (ties >> b) != 0; b++) {
                            double e[NL];
                            load_e(b, e);
                            if ((ties >> b) & 1u) {
                                const int bk = sgb_seq_step_r<NL>(T1s, e, m1r, tr_self, tr_off);
                                back[0] = (b == 0) ? bk : back[0]; back[1] = (b == 1) ? bk : back[1];
                                back[2] = (b == 2) ? bk : back[2]; back[3] = (b == 3) ? bk : back[3];
                            } else { bool t_; sgb_seq_step_fast<NL>(T1s, e, m1r, tr_self, tr_off, t_); }
                        }
                    } else acc += 7;
                }
                __syncwarp();
            }
        }
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int t = 4 * c + b;
            if (FEAT & 2) { stage[(t & 31) * 33 + lane] = (unsigned char)back[b]; lastback = (t == len - 1) ? back[b] : lastback; }
            else acc += back[b];
        }
        if (FEAT & 8) { __syncwarp(); asm volatile("" ::: "memory"); }
    }
    long long t1 = clock64();
    out[threadIdx.x + blockIdx.x * blockDim.x] = m1 + acc + lastback + stage[lane];
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int NPL, int FEAT>
static void run(const char* what, double* d_out, long long* d_cyc, int blocks = 148)
{
    const int n = 1024;
    long long h = 0;
    const size_t smem = 4 * (NPL * 4096 + 1088);
    cudaFuncSetAttribute(k_chunk<NPL, FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int r = 0; r < 2; r++) k_chunk<NPL, FEAT><<<blocks, 128, smem>>>(d_out, d_cyc, n, -0.0513, -5.39, 1 << 30);
    cudaMemcpy(&h, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("planes %d feat %2d %-60s %.1f cycles/step (%s)\n", NPL, FEAT, what, (double)h / (4.0 * n), cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    double* d_out; long long* d_cyc;
    cudaMalloc(&d_out, 8 * 128 * 296); cudaMalloc(&d_cyc, 32);
    run<3, 0>("steps only", d_out, d_cyc);
    run<3, 1>("+ ties/vote (no redo code)", d_out, d_cyc);
    run<3, 5>("+ ties/vote/redo with T1s", d_out, d_cyc);
    run<3, 7>("+ stage/lastback", d_out, d_cyc);
    run<3, 15>("+ syncwarp per chunk", d_out, d_cyc);
    run<3, 15>("same, 2 CTAs per SM", d_out, d_cyc, 296);
    run<3, 31>("everything, two states -Inf", d_out, d_cyc);
    run<1, 31>("everything, two states -Inf", d_out, d_cyc);
    run<1, 0>("steps only", d_out, d_cyc);
    run<1, 5>("+ ties/vote/redo with T1s", d_out, d_cyc);
    run<1, 15>("everything", d_out, d_cyc);
    run<1, 15>("everything, 2 CTAs per SM", d_out, d_cyc, 296);
    run<2, 15>("everything", d_out, d_cyc);
    return 0;
}

// Microbenchmark (development aid): FP64 dependent-issue latency and single-warp throughput on sm_100a.
// One warp per SM sub-partition (128 threads, 1 CTA per SM): the configuration k_mpc_slots runs in.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/_trace/fp64_lat tools/fp64_lat.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, int OP>
__global__ void __launch_bounds__(128) k(double* out, long long* cyc, int iters, double a, double b) {
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                if (OP == 0) x[i] = fma(x[i], a, b);
                if (OP == 1) x[i] = x[i] * a;
                if (OP == 2) x[i] = x[i] + b;
                if (OP == 3) x[i] = 1.0 / x[i] + b;          // IEEE division
                if (OP == 4) x[i] = rsqrt(x[i]) + b;
                if (OP == 5) x[i] = fmin(fmax(x[i] * a, -b), b);   // clip
                if (OP == 6) { double s, c; sincos(x[i], &s, &c); x[i] = s + c; }
                if (OP == 7) x[i] = __drcp_rn(x[i]) + b;
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 12345.678) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int ILP, int OP>
void run(const char* name, double* out, long long* cyc, int threads) {
    int iters = (OP >= 3) ? 64 : 1024;
    k<ILP, OP><<<148, threads>>>(out, cyc, iters, 0.9999999, 1e-7);
    k<ILP, OP><<<148, threads>>>(out, cyc, iters, 0.9999999, 1e-7);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-8s threads/CTA %4d ILP %d : %.2f cycles per op-group (%.2f per op)\n", name, threads, ILP,
           (double)c / (iters * 16.0), (double)c / (iters * 16.0 * ILP));
}

int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 64); cudaMalloc(&cyc, 64);
    for (int threads : {128, 256, 512}) {
        run<1, 0>("DFMA", out, cyc, threads); run<2, 0>("DFMA", out, cyc, threads); run<4, 0>("DFMA", out, cyc, threads);
        run<8, 0>("DFMA", out, cyc, threads);
        run<1, 1>("DMUL", out, cyc, threads); run<1, 2>("DADD", out, cyc, threads);
        run<1, 3>("DDIV", out, cyc, threads); run<4, 3>("DDIV", out, cyc, threads);
        run<1, 7>("DRCP", out, cyc, threads);
        run<1, 4>("RSQRT", out, cyc, threads); run<1, 5>("CLIP", out, cyc, threads);
        run<1, 6>("SINCOS", out, cyc, threads); run<4, 6>("SINCOS", out, cyc, threads);
    }
    return 0;
}

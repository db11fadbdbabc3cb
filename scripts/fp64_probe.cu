// FP64 pipe probe for sm_100a: dependent-chain latency and issue behaviour of DFMA, alone and
// mixed with integer work.   nvcc -arch=sm_100a -O3 -o fp64_probe fp64_probe.cu && ./fp64_probe
#include <cstdio>
#include <cuda_runtime.h>
template <int C, int INTS>
__global__ void probe(double* out, long long* cyc, int iters, double m, double c0) {
    double a[C];
    unsigned x = threadIdx.x * 2654435761u + 1u, y = blockIdx.x + 12345u;
#pragma unroll
    for (int i = 0; i < C; i++) a[i] = 1.0 + i * 0.125 + threadIdx.x * 1e-9;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < C; i++) a[i] = fma(a[i], m, c0);
#pragma unroll
        for (int i = 0; i < INTS; i++) { x = (x << 13) ^ y; y = (y >> 7) + x; }
    }
    const long long t1 = clock64();
    double s = 0; for (int i = 0; i < C; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (double)(x ^ y);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int C, int INTS>
void run(int warps_per_smsp, double* out, long long* cyc) {
    const int iters = 4096, threads = 128 * warps_per_smsp;
    probe<C, INTS><<<148, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
    probe<C, INTS><<<148, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
    long long h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
    const double per_iter = avg / iters;
    printf("warps/SMSP %d chains %d ints/iter %2d : %7.2f cycles/iter  -> %5.2f cycles per DFMA per SMSP, fp64 pipe %5.1f%% (2 cycles/warp-instr)\n",
           warps_per_smsp, C, 2 * INTS, per_iter, per_iter / (C * warps_per_smsp), 100.0 * 2 * C * warps_per_smsp / per_iter);
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(double)); cudaMalloc(&cyc, 148 * sizeof(long long));
    for (int w : {1, 2, 4, 8}) { run<1, 0>(w, out, cyc); run<2, 0>(w, out, cyc); run<4, 0>(w, out, cyc); run<8, 0>(w, out, cyc); }
    for (int w : {4, 8}) { run<2, 1>(w, out, cyc); run<2, 2>(w, out, cyc); run<2, 4>(w, out, cyc); run<4, 2>(w, out, cyc); run<4, 4>(w, out, cyc); run<4, 8>(w, out, cyc); }
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

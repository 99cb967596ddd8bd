// Measures the non-tensor FP64 and FP32 instruction throughput of the device (DADD+DMUL mix, DFMA, FFMA).
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(double *out, int iters) {
    double a0 = threadIdx.x * 1e-9 + 1.0, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; i++) {
        if (MODE == 0) {  // separate mul and add (what a no-FMA bit-exact kernel issues)
#define S(v) v = __dadd_rn(__dmul_rn(v, b), c);
            S(a0) S(a1) S(a2) S(a3) S(a4) S(a5) S(a6) S(a7)
#undef S
        } else {
#define S(v) v = __fma_rn(v, b, c);
            S(a0) S(a1) S(a2) S(a3) S(a4) S(a5) S(a6) S(a7)
#undef S
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
__global__ void kf(float *out, int iters) {
    float a0 = threadIdx.x * 1e-6f + 1.0f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const float b = 1.0000001f, c = 1e-6f;
    for (int i = 0; i < iters; i++) {
#define S(v) v = __fmaf_rn(v, b, c);
        S(a0) S(a1) S(a2) S(a3) S(a4) S(a5) S(a6) S(a7)
#undef S
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 4, threads = 512, iters = 20000;
    double *out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int mode = 0; mode < 3; mode++) {
        float best = 1e30f;
        for (int rep = 0; rep < 5; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<blocks, threads>>>(out, iters);
            else if (mode == 1) k<1><<<blocks, threads>>>(out, iters);
            else kf<<<blocks, threads>>>((float *) out, iters);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double instr = (double) blocks * threads * iters * 8 * (mode == 0 ? 2 : 1);
        printf("%s: %.3f ms, %.2f T thread-instr/s, %.1f instr/clk/SM at %d MHz (%d SMs)\n",
               mode == 0 ? "DMUL+DADD" : (mode == 1 ? "DFMA" : "FFMA"), best, instr / best * 1e-9,
               instr / (best * 1e-3) / p.multiProcessorCount / (p.clockRate * 1e3), p.clockRate / 1000, p.multiProcessorCount);
    }
    return 0;
}

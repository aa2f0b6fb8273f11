// Measures the FP64 vector-pipe peak of the device: (a) dependent-free DFMA, (b) separate DMUL + DADD (what a
// -fmad=false kernel can use).  Prints GFLOP/s counted as 2 flop per DFMA and 1 flop per DMUL / DADD.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k_peak(double *out, int iters, double a, double b) {
    double x[8];
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; it++) {
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) x[i] = __fma_rn(x[i], a, b);
            else { x[i] = __dmul_rn(x[i], a); x[i] = __dadd_rn(x[i], b); }
        }
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int blocks = sms * 8, threads = 256, iters = 20000;
    double *out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k_peak<0><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            else k_peak<1><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            double ops = double(blocks) * threads * iters * 8;     // FP64 instructions per mode-0; x2 for mode 1
            double flops = 2 * ops;
            printf("{\"mode\": \"%s\", \"ms\": %.3f, \"gflops\": %.1f, \"fp64_instr_per_s\": %.4e, \"sms\": %d}\n",
                   mode == 0 ? "dfma" : "dmul+dadd", ms, flops / ms * 1e-6, (mode == 0 ? ops : 2 * ops) / ms * 1e3, sms);
        }
    }
    return 0;
}

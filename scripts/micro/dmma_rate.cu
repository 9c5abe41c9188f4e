// FP64 tensor-core (mma.sync.m8n8k4.f64) issue rate on this GPU vs plain DFMA.  nvcc -arch=sm_100a -o dmma_rate dmma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void dmma_kernel(double* out, int iters) {
    double c[8][2];
    for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = -c[i][0]; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0; for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void dfma_kernel(double* out, int iters) {
    double c[16];
    for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 1e-3 + i;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = fma(c[i], a, b);
    }
    double s = 0; for (int i = 0; i < 16; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* d; cudaMalloc(&d, sizeof(double) * sms * 8 * 256);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int warps = 1; warps <= 8; warps *= 2) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0); dmma_kernel<<<sms * 4, warps * 32>>>(d, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fl = 2.0 * 256 * 8 * (double)iters * warps * sms * 4;
            if (rep) printf("DMMA m8n8k4: %d warps/CTA x4 CTA/SM: %.2f TFLOP/s  (%.1f cycles/mma/warp @1.9GHz)\n", warps, fl / ms / 1e9, ms * 1e-3 * 1.9e9 / (8.0 * iters));
            cudaEventRecord(e0); dfma_kernel<<<sms * 4, warps * 32>>>(d, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
            fl = 2.0 * 32 * 16 * (double)iters * warps * sms * 4;
            if (rep) printf("DFMA       : %d warps/CTA x4 CTA/SM: %.2f TFLOP/s\n", warps, fl / ms / 1e9);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

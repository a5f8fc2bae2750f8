// Issue rate of FFMA vs FFMA2 (fma.rn.f32x2) on sm_100a: both loops keep 8 independent accumulators per thread.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/ffma2_rate tools/micro/ffma2_rate.cu
// Result (B200, 1 block of 128..1024 threads per SM): see the comment at the end of the file.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__global__ void k1(float* out, int iters, float a, float b) {
    float x[8];
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
    float s = 0; for (int i = 0; i < 8; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k2(float* out, int iters, float a, float b) {
    unsigned long long x[8], aa, bb;
    float2 t = make_float2(a, a), u = make_float2(b, b);
    aa = *reinterpret_cast<unsigned long long*>(&t); bb = *reinterpret_cast<unsigned long long*>(&u);
    for (int i = 0; i < 8; ++i) { float2 v = make_float2(threadIdx.x + i, i); x[i] = *reinterpret_cast<unsigned long long*>(&v); }
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = fma2(x[i], aa, bb);
    float s = 0; for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&x[i]); s += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float* out; cudaMalloc(&out, 148 * 1024 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int threads : {128, 512, 1024}) {
        float ms1, ms2;
        k1<<<148, threads>>>(out, 10, 1.0001f, 0.5f); k2<<<148, threads>>>(out, 10, 1.0001f, 0.5f);
        cudaEventRecord(e0); k1<<<148, threads>>>(out, iters, 1.0001f, 0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms1, e0, e1);
        cudaEventRecord(e0); k2<<<148, threads>>>(out, iters, 1.0001f, 0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms2, e0, e1);
        const double inst = (double)iters * 8 * threads / 32;      // warp instructions per SM
        printf("threads %4d: FFMA %.3f ms (%.2f warp-inst/clk/SM @1.9GHz)  FFMA2 %.3f ms (%.2f)\n", threads, ms1,
               inst / (ms1 * 1e-3 * 1.9e9), ms2, inst / (ms2 * 1e-3 * 1.9e9));
    }
    return 0;
}

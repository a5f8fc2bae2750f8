// Micro-probe: what does a small co-resident kernel cost the persistent GEMMs? 64-thread CTAs without shared memory
// (max-shared carveout) that, for `ns` nanoseconds, (0) only sleep, (1) poll a flag with ld.acquire.sys, (2) poll it
// with ld.relaxed.sys, (3) poll with ld.acquire.gpu, (4) stream 16-byte loads/stores over a local buffer.
//   nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o tools/micro/libspin.so tools/micro/spin.cu
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__global__ void __launch_bounds__(64, 16) spin_kernel(int mode, long long ns, unsigned* flag, float* buf, long long nvec, int sleep_ns) {
    const unsigned long long t0 = gtime();
    unsigned acc = 0;
    long long i = (long long)blockIdx.x * 64 + threadIdx.x;
    while (gtime() - t0 < (unsigned long long)ns) {
        if (mode == 1 && threadIdx.x == 0) { unsigned v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); acc += v; }
        if (mode == 2 && threadIdx.x == 0) { unsigned v; asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); acc += v; }
        if (mode == 3 && threadIdx.x == 0) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); acc += v; }
        if (mode == 4) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float* p = buf + 4 * ((i + u * 64LL * gridDim.x) % nvec);
                asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(p) : "memory");
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                float* p = buf + 4 * ((i + u * 64LL * gridDim.x) % nvec);
                asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v[u].x + 1.f), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
            }
            i += 4 * 64LL * gridDim.x;
        }
        if (sleep_ns > 0) __nanosleep(sleep_ns);
    }
    if (acc == 0xFFFFFFFFu) flag[1] = acc;
}
extern "C" int spin_launch(int mode, int ctas, long long ns, unsigned* flag, float* buf, long long nvec, int sleep_ns, void* stream) {
    static bool done = false;
    if (!done) { cudaFuncSetAttribute((const void*)spin_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared); done = true; }
    spin_kernel<<<ctas, 64, 0, (cudaStream_t)stream>>>(mode, ns, flag, buf, nvec, sleep_ns);
    return (int)cudaGetLastError();
}
